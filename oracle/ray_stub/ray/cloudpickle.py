from pickle import *  # noqa: F401,F403
