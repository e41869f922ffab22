"""Drop-in mirror of the reference's `dICP` package (external/dICP/dICP) over the CUDA library."""
from .ICP import ICP  # noqa: F401
from .nn import nn  # noqa: F401
from .loss import loss  # noqa: F401
