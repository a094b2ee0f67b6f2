"""Same import path as the reference: ``from rcnn.config import config``."""
from ..config import config  # noqa: F401
