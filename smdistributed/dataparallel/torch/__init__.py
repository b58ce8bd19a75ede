"""PyTorch binding of the ``smddp`` backend served by b200ddp (see ``torch_smddp.py``)."""
