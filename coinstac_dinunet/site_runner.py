"""``from coinstac_dinunet.site_runner import SiteRunner`` (``comps/fs/site_run.py:1``)."""
from dinunet_implementations_b200.site_runner import SiteRunner  # noqa: F401
