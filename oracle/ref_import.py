"""TEST INFRASTRUCTURE ONLY -- imports the *unmodified* reference modules.

Loads the reference's own PyTorch implementation of the hot path straight from
``/root/reference`` (read-only checkout) under a stub ``tvm`` package, so that
``oracle/gen_golden.py`` can produce golden vectors and ``tests/`` can pin the
restated oracle (``oracle/torch_oracle.py``) against the real thing.

``/root/reference`` exists only in the build container; on the GPU box this
module raises ``ReferenceUnavailable`` and nothing at run time depends on it.

Modules loaded:
  python/tvm/auto_scheduler/dataset.py
  python/tvm/auto_scheduler/cost_model/mlp_model.py
  scripts/pre_experiments/model_myself/utils/{model,select,common}.py
  scripts/vae_experiments/models/{vae,regression}.py
  python/tvm/auto_scheduler/feature.py (unpack_feature only; load_reference_feature())
  scripts/pre_experiments/for_inference/tuning_improved.py (select_features only; load_reference_tuning())
"""
import importlib.machinery
import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("VAEX_REFERENCE_ROOT", "/root/reference")


class ReferenceUnavailable(RuntimeError):
    pass


def available():
    return os.path.isfile(os.path.join(
        REF_ROOT, "python/tvm/auto_scheduler/cost_model/mlp_model.py"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, None)   # torch._dynamo calls find_spec on every loaded module
    m.__dict__.update(attrs)
    m.__path__ = []  # behave like a package
    sys.modules[name] = m
    return m


def _load(modname, relpath):
    path = os.path.join(REF_ROOT, relpath)
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    return mod


_CACHE = {}


def load_reference():
    """Returns a namespace with the reference modules: .mlp_model, .dataset,
    .vae_model (utils/model.py), .select, .common, .seg_vae, .seg_reg."""
    if "ns" in _CACHE:
        return _CACHE["ns"]
    if not available():
        raise ReferenceUnavailable("%s not present (GPU box?)" % REF_ROOT)
    import numpy as np

    if "tvm" in sys.modules and not getattr(sys.modules["tvm"], "_vaex_stub", False):
        raise RuntimeError("a real tvm is already imported")

    class _Dummy:  # MeasureInput / MeasureResult / RecordReader placeholders
        def __init__(self, *a, **k):
            pass

    def _nope(*a, **k):
        raise NotImplementedError("TVM C++ feature extraction is not available in the stub")

    tvm = _stub("tvm", _vaex_stub=True)
    asch = _stub("tvm.auto_scheduler")
    _stub("tvm.auto_scheduler.measure_record", RecordReader=_Dummy)
    _stub("tvm.auto_scheduler.measure", MeasureInput=_Dummy, MeasureResult=_Dummy)
    _stub("tvm.auto_scheduler.feature",
          get_per_store_features_from_measure_pairs=_nope,
          get_per_store_features_from_states=_nope,
          get_per_store_features_from_file=_nope)
    cm = _stub("tvm.auto_scheduler.cost_model")
    _stub("tvm.auto_scheduler.cost_model.xgb_model",
          get_workload_embedding=lambda key: np.zeros(9, dtype=np.float32))

    class PythonBasedModel:
        def __init__(self):
            pass

    _stub("tvm.auto_scheduler.cost_model.cost_model", PythonBasedModel=PythonBasedModel)
    tvm.auto_scheduler = asch
    asch.cost_model = cm

    ns = types.SimpleNamespace()
    ns.dataset = _load("tvm.auto_scheduler.dataset", "python/tvm/auto_scheduler/dataset.py")
    ns.mlp_model = _load("tvm.auto_scheduler.cost_model.mlp_model",
                         "python/tvm/auto_scheduler/cost_model/mlp_model.py")
    base = "scripts/pre_experiments/model_myself/utils/"
    ns.vae_model = _load("_vaex_ref_utils_model", base + "model.py")
    ns.select = _load("_vaex_ref_utils_select", base + "select.py")
    ns.common = _load("_vaex_ref_utils_common", base + "common.py")
    ns.seg_vae = _load("_vaex_ref_seg_vae", "scripts/vae_experiments/models/vae.py")
    try:
        ns.seg_reg = _load("_vaex_ref_seg_reg", "scripts/vae_experiments/models/regression.py")
    except Exception as e:  # regression.py may import its sibling by package path
        ns.seg_reg = None
        ns.seg_reg_error = repr(e)
    _CACHE["ns"] = ns
    return ns


def load_reference_feature():
    """The reference's python/tvm/auto_scheduler/feature.py (for unpack_feature, :49-158) under the stub
    package; its C++ entry points (_ffi_api) are placeholders that raise."""
    if "feature" in _CACHE:
        return _CACHE["feature"]
    load_reference()

    class _Dummy:
        pass

    _stub("tvm.auto_scheduler.loop_state", State=_Dummy, StateObject=_Dummy)
    _stub("tvm.auto_scheduler._ffi_api")
    mod = _load("tvm.auto_scheduler._ref_feature", "python/tvm/auto_scheduler/feature.py")
    _CACHE["feature"] = mod
    return mod


def load_reference_training():
    """The reference's utils/training.py (losses + compute_total_loss, :133-226) with `utils.model` / `utils.common`
    bound to the already loaded reference modules."""
    if "training" in _CACHE:
        return _CACHE["training"]
    ns = load_reference()
    utils = _stub("utils")
    sys.modules["utils.model"] = ns.vae_model
    sys.modules["utils.common"] = ns.common
    utils.model, utils.common = ns.vae_model, ns.common
    mod = _load("_vaex_ref_utils_training", "scripts/pre_experiments/model_myself/utils/training.py")
    _CACHE["training"] = mod
    return mod


def load_reference_extent():
    """The reference's utils/extent.py (state_to_records and the schedule clean-up, :8-136) under the stub tvm."""
    if "extent" in _CACHE:
        return _CACHE["extent"]
    load_reference()
    mod = _load("_vaex_ref_utils_extent", "scripts/pre_experiments/model_myself/utils/extent.py")
    _CACHE["extent"] = mod
    return mod


def load_reference_seg_trainer():
    """The reference's scripts/vae_experiments/trainer.py (VAE_Trainer.vae_loss :270-280, train_vae :33-118) with its
    `models` package bound to the reference directory and `util_manager`'s TVM imports stubbed."""
    if "seg_trainer" in _CACHE:
        return _CACHE["seg_trainer"]
    ns = load_reference()
    base = os.path.join(REF_ROOT, "scripts/vae_experiments")
    _stub("tvm.relay", testing=None)
    init = _stub("tvm.relay.testing.init", create_workload=lambda *a, **k: None)
    testing = _stub("tvm.relay.testing", init=init)
    sys.modules["tvm.relay"].testing = testing
    sys.modules["tvm"].relay = sys.modules["tvm.relay"]
    models = _stub("models")
    models.__path__ = [os.path.join(base, "models")]
    sys.modules["models.vae"] = ns.seg_vae
    if ns.seg_reg is not None:
        sys.modules["models.regression"] = ns.seg_reg
    _load("util_manager", "scripts/vae_experiments/util_manager.py")
    mod = _load("_vaex_ref_seg_trainer", "scripts/vae_experiments/trainer.py")
    _CACHE["seg_trainer"] = mod
    return mod


def load_reference_tuning():
    """The reference's for_inference/tuning_improved.py (select_features, :11-175) under the stub tvm package."""
    if "tuning" in _CACHE:
        return _CACHE["tuning"]
    load_reference()
    mod = _load("_vaex_ref_tuning_improved", "scripts/pre_experiments/for_inference/tuning_improved.py")
    _CACHE["tuning"] = mod
    return mod
