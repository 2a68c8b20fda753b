from dinunet_implementations_b200.metrics import *  # noqa: F401,F403
