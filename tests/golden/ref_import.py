"""Import the *unmodified* reference from /root/reference on CPU (dev container only).

The reference cannot travel to the GPU box, so this module is used by exactly two
things: ``make_golden.py`` (which writes the committed ``*.npz`` fixtures) and the
optional live cross-check in ``tests/test_oracle_vs_reference_live.py`` (skipped
when /root/reference is absent).  Nothing in the product imports it.

Three runtime shims repair library drift unrelated to the hot-path arithmetic
(SURVEY.md section 8(c)); no reference file is edited or copied:
  1. scipy.signal.gaussian          (models/net_canny.py:4)
  2. a stub ``visdom`` module       (cal_prototype.py:12, segmentation_evaluate.py:11)
  3. torchvision.models.utils       (models/thermal_semantic_segmentation_models.py:9)
"""
import os
import sys
import types

REF_ROOT = os.environ.get("TSS_REFERENCE_ROOT", "/root/reference")


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
    for name in ("utils.eval_tools", "utils.loss", "cal_prototype", "self_training",
                 "generate_pseudo_label", "data.freiburg_dataset"):
        ns.__dict__[name.replace(".", "_")] = importlib.import_module(name)
    return ns
