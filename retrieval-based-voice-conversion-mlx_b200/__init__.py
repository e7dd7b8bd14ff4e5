"""rvc_b200 - B200-native NSF-HiFiGAN decoder (drop-in for the reference's HiFiGANNSFGenerator).

Only the decoder hot path lives here; see DESIGN.md for scope.  Importing this package does not
load the CUDA library; the first ``forward`` (or ``_lib.lib()``) does, and fails loudly if it is
missing.
"""
from .decoder import HiFiGANGenerator, HiFiGANMRFGenerator, HiFiGANNSFGenerator, ResBlock, SourceModuleHnNSF  # noqa: F401
from . import _lib  # noqa: F401
from . import chunking  # noqa: F401
from . import configs  # noqa: F401
from . import loader  # noqa: F401
from .loader import load_checkpoint  # noqa: F401
from .synthesizer import Synthesizer  # noqa: F401
from .refinegan import RefineGANGenerator  # noqa: F401
from .rmvpe import RMVPE0Predictor  # noqa: F401
from . import rmvpe  # noqa: F401
from .hubert import HubertModelWithFinalProj  # noqa: F401
from . import hubert  # noqa: F401
from . import runner  # noqa: F401
from .runner import MultiGpuDecoder  # noqa: F401

__all__ = ["HiFiGANGenerator", "HiFiGANMRFGenerator", "HiFiGANNSFGenerator", "ResBlock", "SourceModuleHnNSF", "chunking", "configs", "loader",
           "load_checkpoint", "runner", "MultiGpuDecoder", "Synthesizer", "RefineGANGenerator", "RMVPE0Predictor", "rmvpe", "HubertModelWithFinalProj", "hubert"]
