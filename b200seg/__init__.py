"""b200seg -- importable name of the package that lives in ``thermal-semantic-segmentation_b200/``.

The directory carries the reference's name (with hyphens, so it cannot be imported directly); this
shim puts it on the package path, so ``b200seg.ops``, ``b200seg.dropin`` ... resolve to the modules
there.  Nothing is implemented here.
"""
import os as _os

_IMPL_DIR = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                          "thermal-semantic-segmentation_b200")
if not _os.path.isdir(_IMPL_DIR):  # pragma: no cover
    raise ImportError("b200seg: implementation directory missing: %s" % _IMPL_DIR)
__path__.append(_IMPL_DIR)

from ._lib import lib, library_path, TssError  # noqa: E402,F401
