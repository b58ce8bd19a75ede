"""b200ddp — import name of the B200-native data-parallel training framework.

The source tree lives in ``cloud-security-pytorch-sagemaker-distributed-workshop_b200/``
(the directory name the build asks for; it is not a legal Python identifier), so this
shim splices that directory into ``b200ddp.__path__`` and re-exports the public API.
"""
import os as _os

_SRC = _os.path.join(
    _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
    "cloud-security-pytorch-sagemaker-distributed-workshop_b200",
)
__path__.insert(0, _SRC)  # b200ddp.parallel, b200ddp.ops, ... resolve inside _SRC

from ._api import *  # noqa: F401,F403,E402
from ._api import __all__, __version__  # noqa: F401,E402
