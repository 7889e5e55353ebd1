from .multi_modal_discussion_transformer import *  # noqa: F401,F403  (mirrors mDT/src/models/__init__.py)
