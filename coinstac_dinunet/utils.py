"""``from coinstac_dinunet.utils import duration`` (``local.py:5``, ``remote.py:5``)."""
from dinunet_implementations_b200.utils import *  # noqa: F401,F403
from dinunet_implementations_b200.utils import duration, set_seed, logger  # noqa: F401
