"""b200ddp's stand-in for ``smdistributed.dataparallel`` (see ``smdistributed/__init__.py``)."""
__version__ = "b200ddp"
