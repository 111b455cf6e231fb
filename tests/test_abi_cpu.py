"""CPU-side checks of the drop-in boundary: the shared library loads without a GPU, exports every symbol
include/b200ppo.h declares, reads the reference's config file verbatim, and refuses to run without a
device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def capi():
    import __graft_entry__ as g
    g.build(oracle=False)
    from reinforcement_learning_cpp_parallel_b200 import capi
    return capi


def test_header_symbols_all_exported(capi):
    hdr = open(os.path.join(ROOT, "include", "b200ppo.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 40
    L = capi.lib()
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in include/b200ppo.h but not exported"
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    assert L.b2_abi_version() == 2


def test_struct_layouts_match_header(capi):
    # sizes the C compiler gives the header's structs (compiled here with gcc)
    import subprocess, tempfile
    src = ('#include "b200ppo.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu\\n", sizeof(b2_config),'
           'sizeof(b2_batch), sizeof(b2_batch_out), sizeof(b2_update_stats));return 0;}')
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(src)
        subprocess.check_call(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), os.path.join(d, "t.c"), "-o",
                               os.path.join(d, "t")])
        sizes = list(map(int, subprocess.check_output([os.path.join(d, "t")]).split()))
    assert sizes == [C.sizeof(capi.Config), C.sizeof(capi.Batch), C.sizeof(capi.BatchOut), C.sizeof(capi.UpdateStats)]


def test_defaults_are_the_reference_hyperparameters(capi):
    c = capi.default_config()
    # config:1-6
    assert (c.control_vars, c.obs_vars, c.n_episodes, c.n_timestep, c.max_timestep, c.update_timestep) == \
        (1, 6, 10000, 100, 1000000, 2000)
    # src/main.cpp:196-203
    assert (c.action_std, c.k_epochs, c.eps_clip, c.gamma, c.lr, c.beta1, c.beta2) == (0.5, 80, 0.2, 0.99, 0.0003, 0.9, 0.999)
    assert (c.hidden1, c.hidden2, c.n_minibatches) == (64, 32, 1)


def test_config_file_format(capi, tmp_path):
    p = tmp_path / "config"
    p.write_text("1 = control_vars\n6 = obs_vars\n500  = Nepisodes \n   50 = Maximum time step in one episode \n"
                 "12345 = Maximum training timestep (either define this or define Nepisodes) \n700 = updateTimestep\n")
    c = capi.read_config(str(p))
    assert (c.control_vars, c.obs_vars, c.n_episodes, c.n_timestep, c.max_timestep, c.update_timestep) == \
        (1, 6, 500, 50, 12345, 700)
    assert c.k_epochs == 80  # untouched fields keep their defaults
    bad = tmp_path / "bad"
    bad.write_text("1 = a\nsix = b\n")
    with pytest.raises(capi.B2Error) as ei:
        capi.read_config(str(bad))
    assert ei.value.code == -5 and "bad:2" in str(ei.value)
    short = tmp_path / "short"
    short.write_text("1\n6\n")
    with pytest.raises(capi.B2Error):
        capi.read_config(str(short))
    with pytest.raises(capi.B2Error):
        capi.read_config(str(tmp_path / "missing"))


def test_no_cpu_fallback(capi):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.B2Error) as ei:
        capi.Engine(capi.default_config())
    assert ei.value.code == -2


def test_argument_validation_happens_before_device_use(capi):
    for kw in (dict(control_vars=2), dict(obs_vars=5), dict(hidden1=100, hidden2=256), dict(hidden1=512, hidden2=64), dict(n_envs=0),
               dict(precision=7), dict(k_epochs=0), dict(action_std=0.0), dict(env_flags=8),
               dict(env_flags=capi.ENV_SWINGUP, precision=capi.PREC_F32)):
        with pytest.raises(capi.B2Error) as ei:
            capi.Engine(capi.default_config(**kw))
        assert ei.value.code == -1, kw


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under the package or include/ may reference it."""
    for base in ("reinforcement_learning_cpp_parallel_b200", "include"):
        for dp, dn, fn in os.walk(os.path.join(ROOT, base)):
            if "build" in dp or "__pycache__" in dp:
                continue
            for f in fn:
                if f.endswith((".so", ".o", ".pyc")):
                    continue
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "liboracle" not in txt and "oracle.h" not in txt and "import oracle" not in txt, (dp, f)
                assert "orc_" not in txt.replace("force", ""), (dp, f)


def test_env_switches_in_design_exist_in_the_sources():
    """every B2_* environment switch DESIGN.md s7b documents is read somewhere in the library (or bench / tools), and
    every getenv("B2_...") of the library is documented there -- the A/B numbers of the document must stay re-measurable"""
    import glob
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    design = open(os.path.join(root, "DESIGN.md")).read()
    sec = design[design.index("## 7b."):design.index("## 7c.")]
    documented = set(re.findall(r"`(B2_[A-Z0-9_]+)", sec))
    src = ""
    for pat in ("reinforcement_learning_cpp_parallel_b200/csrc/*.cu", "reinforcement_learning_cpp_parallel_b200/csrc/*.cuh",
                "reinforcement_learning_cpp_parallel_b200/csrc/*.h", "reinforcement_learning_cpp_parallel_b200/csrc/*.cpp"):
        for f in glob.glob(os.path.join(root, pat)):
            src += open(f).read()
    read = set(re.findall(r'getenv\("(B2_[A-Z0-9_]+)"\)', src))
    assert documented <= read, sorted(documented - read)
    diagnostics = {"B2_WIDE_SKIP", "B2_FH_DBG", "B2_GEMM_DBG"}   # profiling aids described next to their code, not switches of the product
    assert read - diagnostics <= documented, sorted(read - diagnostics - documented)
