"""Native ops: ctypes binding (``_lib``), autograd wrappers (``functional``), build script."""
from . import _lib  # noqa: F401


def native_available() -> bool:
    return _lib.available()
