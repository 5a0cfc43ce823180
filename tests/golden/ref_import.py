"""Import the *unmodified* reference: from /root/reference (dev container) or, on the GPU box, from the
git-ignored copy ``baseline/_ref/`` that ``__graft_entry__.build()`` stages there.

Used by ``make_golden.py`` (which writes the committed ``*.npz`` fixtures), by the live cross-check
``tests/test_oracle_vs_reference_live.py``, by ``tests/test_gpu_reference_dropin.py`` (the reference's own modules
with ``dropin.install()`` applied) and by ``bench.py --impl reference``.  Nothing in the product imports it.

Three runtime shims repair library drift unrelated to the hot-path arithmetic
(SURVEY.md section 8(c)); no reference file is edited or copied:
  1. scipy.signal.gaussian          (models/net_canny.py:4)
  2. a stub ``visdom`` module       (cal_prototype.py:12, segmentation_evaluate.py:11)
  3. torchvision.models.utils       (models/thermal_semantic_segmentation_models.py:9)
"""
import os
import sys
import types

_REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _default_root() -> str:
    for cand in (os.environ.get("TSS_REFERENCE_ROOT"), "/root/reference", os.path.join(_REPO, "baseline", "_ref")):
        if cand and os.path.isfile(os.path.join(cand, "self_training.py")):
            return cand
    return "/root/reference"


REF_ROOT = _default_root()


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "self_training.py"))


def import_reference():
    """Returns a namespace with the reference modules of the hot path."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    import torch  # noqa: F401
    import scipy.signal
    import scipy.signal.windows
    if not hasattr(scipy.signal, "gaussian"):
        scipy.signal.gaussian = scipy.signal.windows.gaussian
    if "visdom" not in sys.modules:
        vis = types.ModuleType("visdom")
        vis.Visdom = lambda *a, **k: None
        sys.modules["visdom"] = vis
    if "torchvision.models.utils" not in sys.modules:
        import torch.hub
        tvu = types.ModuleType("torchvision.models.utils")
        tvu.load_state_dict_from_url = torch.hub.load_state_dict_from_url
        sys.modules["torchvision.models.utils"] = tvu
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    ns = types.SimpleNamespace()
    import importlib
    for name in ("utils.eval_tools", "utils.loss", "utils", "data", "cal_prototype", "self_training",
                 "generate_pseudo_label", "data.freiburg_dataset", "segmentation_evaluate"):
        ns.__dict__[name.replace(".", "_")] = importlib.import_module(name)
    return ns
