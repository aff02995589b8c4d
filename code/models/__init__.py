"""`models` package of the DeepGraft code tree (code/models/__init__.py:3 in the
reference exports TransMIL).  Only the hot-path plugin lives here."""
from .TransMIL import TransMIL  # noqa: F401
