"""Drop-in namespace for the absent ``coinstac-dinunet`` package.

The reference's files import ``COINNLocal`` / ``COINNRemote`` (``local.py:4``,
``remote.py:4``), ``COINNDataset`` / ``COINNTrainer`` / ``COINNDataHandle``
(``comps/fs/__init__.py:6``, ``comps/icalstm/__init__.py:7``), ``utils.duration``
(``local.py:5``) and ``site_runner.SiteRunner`` (``comps/fs/site_run.py:1``) from here.
Everything is implemented in ``dinunet_implementations_b200``; this package only re-exports
it so those files run unmodified.
"""
from dinunet_implementations_b200 import (  # noqa: F401
    COINNLocal, COINNRemote, COINNTrainer, NNTrainer, COINNDataset, COINNDataHandle, SiteRunner,
)
from dinunet_implementations_b200 import metrics, distrib  # noqa: F401
from . import utils, site_runner  # noqa: F401
