"""apvae — host-side mirror of the reference's evaluation / sampling call surface
(experiment/dmm/dmm_*.py builders, tfsnippet Evaluator / VariationalInference)
over the sm_100a CUDA library libapvae_b200.so.  PyTorch is used only for device
buffers, streams and torch.distributed; all arithmetic of the path runs in the
library's own kernels.  No CPU fallback."""
from . import _lib                                   # noqa: F401  (raises if the .so is missing)
from ._lib import ApvaeError                         # noqa: F401
from .plan import Plan, conv2d, deconv2d, dense      # noqa: F401
