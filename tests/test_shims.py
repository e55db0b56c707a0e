"""The reference's drivers against ``shims/``: with that directory first on ``sys.path`` the UNMODIFIED
``02_run_sae_baseline.py``, ``run_generation.py`` and ``01_reproduce_logit_lens.py`` import (every bare module
they name resolves: models, feature_map, metrics, utils, plots, run_generation, sae_lens), and every function the
shims keep has the reference's parameter list (src/models.py:8,55,97,173; run_generation.py:15,21,32;
plots.py:4; utils.py:5; metrics.py). Runs only where /root/reference exists (not on the GPU box).

Done in a subprocess so that bare module names like ``metrics`` do not leak into this test session."""
import ast
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = "/root/reference/src"

pytestmark = pytest.mark.skipif(not os.path.isdir(REF_SRC), reason="the reference tree is not on this machine")

CHILD = r"""
import importlib.util, inspect, json, sys, types
root, ref = sys.argv[1], sys.argv[2]
sys.path[:0] = [root + "/shims", root]
try:
    import matplotlib.pyplot  # noqa: F401
except Exception:               # 01 imports pyplot at module level; a stub is enough to import it
    m, p = types.ModuleType("matplotlib"), types.ModuleType("matplotlib.pyplot")
    m.pyplot = p
    sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = m, p
out = {"imported": {}, "signatures": {}, "origins": {}}
for name, fn in (("ref02", "02_run_sae_baseline.py"), ("refgen", "run_generation.py"), ("ref01", "01_reproduce_logit_lens.py")):
    spec = importlib.util.spec_from_file_location(name, ref + "/" + fn)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out["imported"][fn] = sorted(k for k in vars(mod) if not k.startswith("__"))
    for k, v in vars(mod).items():
        origin = getattr(v, "__module__", None)
        if origin and origin.startswith("taboo_brittleness_b200"):
            out["origins"][fn + ":" + k] = origin
def sig(f):
    return [[p.name, repr(p.default) if p.default is not inspect.Parameter.empty else None, p.kind.name]
            for p in inspect.signature(f).parameters.values()]
import models, metrics, utils, plots, run_generation, feature_map, sae_lens
for modname, mod in (("models", models), ("metrics", metrics), ("utils", utils), ("plots", plots), ("run_generation", run_generation)):
    for k, v in vars(mod).items():
        if inspect.isfunction(v):
            out["signatures"][modname + "." + k] = sig(v)
out["feature_map_words"] = sorted(feature_map.feature_map)
out["sae_has"] = [hasattr(sae_lens.SAE, "from_pretrained"), hasattr(sae_lens.SAE, "encode")]
out["from_pretrained"] = sig(sae_lens.SAE.from_pretrained)
print("RESULT " + json.dumps(out))
"""


def _ref_signatures(path):
    """{function name: [[param, default repr or None, kind], ...]} of the top-level functions of a reference
    file, from its syntax tree (the file itself cannot be imported: nnsight / sae_lens are missing)."""
    tree = ast.parse(open(path).read())
    sigs = {}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef):
            a = node.args
            names = [x.arg for x in a.posonlyargs + a.args]
            defaults = [None] * (len(names) - len(a.defaults)) + [repr(ast.literal_eval(dflt)) for dflt in a.defaults]
            sigs[node.name] = [[n, dflt, "POSITIONAL_OR_KEYWORD"] for n, dflt in zip(names, defaults)]
    return sigs


@pytest.fixture(scope="module")
def child():
    env = dict(os.environ, PYTHONPATH="")
    proc = subprocess.run([sys.executable, "-c", CHILD, ROOT, REF_SRC], capture_output=True, text=True, env=env,
                          timeout=600, cwd=ROOT)
    assert proc.returncode == 0, proc.stderr[-3000:]
    line = [ln for ln in proc.stdout.splitlines() if ln.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


def test_unmodified_drivers_import_against_the_shims(child):
    got = child["imported"]
    assert {"get_top_sae_features", "latents_to_word_guesses", "analyze_sae_baseline", "SAE", "feature_map",
            "calculate_metrics", "find_model_response_start"} <= set(got["02_run_sae_baseline.py"])
    assert {"generate_for_word", "setup_model", "get_model_response", "get_layer_logits",
            "clean_gpu_memory"} <= set(got["run_generation.py"])
    assert {"aggregate_response_logits", "evaluate_single_word", "plot_token_probability", "calculate_metrics",
            "clean_gpu_memory", "setup_model"} <= set(got["01_reproduce_logit_lens.py"])
    # ... and what they imported is this package, not something else of the same name
    o = child["origins"]
    assert o["02_run_sae_baseline.py:SAE"] == "taboo_brittleness_b200.sae"
    assert o["02_run_sae_baseline.py:find_model_response_start"] == "taboo_brittleness_b200.models"
    assert o["run_generation.py:get_layer_logits"] == "taboo_brittleness_b200.models"
    assert o["01_reproduce_logit_lens.py:calculate_metrics"] == "taboo_brittleness_b200.metrics"
    assert child["sae_has"] == [True, True]
    assert [p[0] for p in child["from_pretrained"]] == ["release", "sae_id", "device"]


@pytest.mark.parametrize("module,names", [
    ("models", ["setup_model", "get_model_response", "get_layer_logits", "find_model_response_start"]),
    ("run_generation", ["load_config", "_pair_paths", "save_pair"]),
    ("utils", ["clean_gpu_memory"]),
    ("plots", ["plot_token_probability"]),
    ("metrics", ["prompt_accuracy_at_k", "any_pass_at_k", "global_majority_vote_at_k", "calculate_metrics"]),
])
def test_kept_functions_have_the_reference_parameter_lists(child, module, names):
    ref = _ref_signatures(os.path.join(REF_SRC, module + ".py"))
    for n in names:
        assert child["signatures"][f"{module}.{n}"] == ref[n], f"{module}.{n}"


def test_feature_map_words_equal_the_reference(child):
    sys.path.insert(0, REF_SRC)
    try:
        import importlib.util
        spec = importlib.util.spec_from_file_location("ref_feature_map", os.path.join(REF_SRC, "feature_map.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path.remove(REF_SRC)
    assert child["feature_map_words"] == sorted(mod.feature_map)
