"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/magic_b200.h declares,
the ctypes binding covers exactly that set, the header is valid C, and -- without a GPU -- the product path fails
loudly instead of falling back."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "magic_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(magic_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    from magic_b200 import _lib, build
    build.build(verbose=False)
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 28
    for n in names:
        assert hasattr(lib, n), "libmagic_b200.so does not export " + n
    assert sorted(_lib.SIGNATURES) == names, (set(names) ^ set(_lib.SIGNATURES))
    assert lib.magic_version() >= 100


def test_header_is_plain_c():
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", HEADER], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_config_struct_layout_matches_header():
    """ctypes mirror vs the C compiler's layout of magic_config / magic_stats."""
    from magic_b200 import _lib
    prog = r'''
    #include <stdio.h>
    #include <stddef.h>
    #include "magic_b200.h"
    int main(void) {
      printf("%zu %zu %zu %zu %zu %zu\n", sizeof(magic_config), offsetof(magic_config, L), offsetof(magic_config, def_weight),
             offsetof(magic_config, seed), sizeof(magic_stats), offsetof(magic_stats, hbm_bytes));
      return 0;
    }'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        src, exe = os.path.join(d, "t.c"), os.path.join(d, "t")
        open(src, "w").write(prog)
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), src, "-o", exe], check=True)
        got = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    cfg, st = _lib.MagicConfig, _lib.MagicStats
    assert got == [C.sizeof(cfg), cfg.L.offset, cfg.def_weight.offset, cfg.seed.offset, C.sizeof(st), st.hbm_bytes.offset]


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for the GPU-less container")
    from magic_b200 import _lib
    from magic_b200.engine import Engine, EngineConfig
    with pytest.raises(_lib.MagicError):
        Engine(EngineConfig(D=1600, encoder_neurons=(32,), latent=4, batch=8))
    lib = _lib.load()
    cfg = EngineConfig(D=1600, encoder_neurons=(32,), latent=4, batch=8).to_c()
    h = C.c_void_p()
    assert lib.magic_create(C.byref(cfg), None, C.byref(h)) != 0
    assert b"no CUDA device" in lib.magic_last_error() or b"CUDA" in lib.magic_last_error()
    assert not h.value


def test_product_code_never_imports_the_oracle():
    for base, _, files in os.walk(os.path.join(ROOT, "magic_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert "oracle" not in open(os.path.join(base, f)).read().replace("the oracle", ""), f
    for base, _, files in os.walk(os.path.join(ROOT, "experiments")):
        for f in files:
            if f.endswith(".py"):
                assert "import oracle" not in open(os.path.join(base, f)).read() and "from oracle" not in open(os.path.join(base, f)).read()


def test_experiment_directories_expose_the_reference_surface():
    import importlib.util
    for exp in ("8c", "8d", "8e", "9a", "11a"):
        p = os.path.join(ROOT, "experiments", exp, "constants.py")
        spec = importlib.util.spec_from_file_location("constants_" + exp, p)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        F = m.FLAGS
        assert F.IMAGE_SIZE == 200 and F.NUM_LAYERS == 3 and F.BACKEND == "b200"
        assert F.BATCH_SIZE == {"8c": 100, "8d": 100, "8e": 100, "9a": 24, "11a": 24}[exp]      # <exp>/constants.py:52
        want = m.DISTANCE_METRICS.SQUARED_DIFFERENCE if exp == "11a" else m.DISTANCE_METRICS.COSINE_DISTANCE
        assert F.DISTANCE_METRIC == want                                                           # <exp>/constants.py:60


REFERENCE_FLAGS = (   # every attribute of 8c/constants.py:12-89 (11a adds RUN_INTERMEDIATE_TESTS, 11a/constants.py:88)
    "DATA_DIR DATA_URL INTERNET ALLOW_DOWNLOADS IMAGE_SIZE NUM_EXAMPLES_TO_LOAD_INTO_QUEUE NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN "
    "NUM_EXAMPLES_PER_EPOCH_FOR_EVAL MAX_EPOCHS TRAINING_STEP MAX_TRAIN_ITERATIONS ROTATE NUM_LAYERS BATCH_SIZE DISTANCE_METRIC "
    "USE_FP16 TOWER_NAME NUM_THREADS NOTIFICATION_EMAIL EMAIL_INFO RESTORE RESTORE_FROM CHECKPOINT_DIR LEARNING_RATE CHECK_DATASET "
    "TRAIN_DIR SUMMARIES_DIR USE_TENSORBOARD KEEP_PROBABILITY RUN_NAME PRINT_INFO SEND_LOSS_TO_POOR_MANS_TENSORBOARD_EVERY_N_STEPS "
    "SAVE_NETSTATE_AND_EMAIL_STATUS_EVERY_N_STEPS").split()


def test_experiment_main_modules_have_the_reference_layout():
    """main.py of every experiment directory: module-level `images_batch, labels_batch = overlap_input.inputs(...)`,
    `n_samples`, `TrainingException`, and `train(sess, batch_size=FLAGS.BATCH_SIZE, training_epochs=60, display_step=5)`
    (8c/main.py:25-33; 11a/main.py:37-42 has no display_step); FLAGS carries every reference attribute."""
    import importlib
    import inspect
    sys.path.insert(0, ROOT)
    for exp in ("8c", "8d", "8e", "9a", "11a"):
        main = importlib.import_module("experiments.%s.main" % exp)
        F = main.FLAGS
        for name in REFERENCE_FLAGS + (["RUN_INTERMEDIATE_TESTS"] if exp == "11a" else []):
            assert hasattr(F, name), (exp, name)
        assert F.NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN == (9600 if exp in ("8c", "8d", "8e") else 960)
        assert main.n_samples == F.NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN
        assert issubclass(main.TrainingException, Exception)
        assert hasattr(main.images_batch, "eval") and hasattr(main.labels_batch, "eval")
        assert main.images_batch.pipeline is main.labels_batch.pipeline
        params = list(inspect.signature(main.train).parameters.values())
        want = ["sess", "batch_size", "training_epochs"] + (["display_step"] if exp != "11a" else [])
        assert [p.name for p in params[:len(want)]] == want, exp
        assert params[1].default == F.BATCH_SIZE and params[2].default == 60
        if exp != "11a":
            assert params[3].default == 5
        oi = importlib.import_module("experiments.%s.overlap_input" % exp)
        assert list(inspect.signature(oi.inputs).parameters) == ["normalize", "reshape", "rotation"]   # 8c/overlap_input.py:76
        if main.images_batch.pipeline.model is None:   # (an earlier test of the same process may have built a TangoEncoder)
            with pytest.raises(RuntimeError):
                main.images_batch.eval()           # no TangoEncoder yet: the handles evaluate on its engine


def test_inputs_requires_a_data_dir():
    from magic_b200 import overlap_input

    class F:
        DATA_DIR = ""
    with pytest.raises(ValueError, match="Please supply a data_dir"):          # 8c/overlap_input.py:98-99
        overlap_input.inputs(normalize=True, reshape=True, rotation=True, flags=F)
