"""``smdistributed`` import name served by b200ddp, so that SageMaker data-parallel scripts run unmodified on this framework.

The real ``smdistributed.dataparallel`` (SMDDP v1.4.1 in the reference's job log,
``notebooks/2_pytorch_dist_smddp_gpu.ipynb:287``) is closed source and not installable offline.  The reference reaches it
through exactly two lines: ``import smdistributed.dataparallel.torch.torch_smddp`` (registers the ``'smddp'`` backend) and
``dist.init_process_group(backend='smddp')`` (``notebooks/code/cifar10-distributed-smddp-gpu.py:17,23``).  Both are provided
here; see ``smdistributed/dataparallel/torch/torch_smddp.py``.
"""
