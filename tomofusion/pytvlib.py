from tomo_tv_b200.pytvlib import *  # noqa: F401,F403
