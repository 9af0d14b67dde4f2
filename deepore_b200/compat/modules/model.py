"""`modules.model` -- the name under which the reference pickles its Model (src/modules/model.py:59)."""
from deepore_b200.model import Model  # noqa: F401
