"""B200-native decentralized (multi-site) neural-network training.

Same capabilities and the same COINSTAC computation API as
``trendscenter/dinunet_implementations`` + its absent engine package ``coinstac_dinunet``
(see SURVEY.md), rebuilt Blackwell-first: one process per GPU, every GPU a site, the
aggregation step and the model hot ops as hand-written sm_100a kernels
(``csrc/``, loaded through ``ops.native``).

Public names (what the reference's ``local.py`` / ``remote.py`` / ``comps/*`` import from
``coinstac_dinunet``; the ``coinstac_dinunet`` shim package re-exports them):
"""
from .config import AggEngine, Phase, Mode
from .data import COINNDataset, COINNDataHandle
from .nn import COINNTrainer, NNTrainer
from .nodes import COINNLocal, COINNRemote
from .site_runner import SiteRunner
from . import utils, metrics, distrib

__version__ = "0.1.0"

__all__ = [
    "AggEngine", "Phase", "Mode", "COINNDataset", "COINNDataHandle", "COINNTrainer", "NNTrainer",
    "COINNLocal", "COINNRemote", "SiteRunner", "utils", "metrics", "distrib",
]
