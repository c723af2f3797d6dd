"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol the header declares
(no compute call without a GPU), and the product fails loudly instead of falling back to the CPU."""
import ctypes
import os

import pytest
import torch

from gcnn_keras_b200 import _lib


def test_library_exists_and_loads():
    assert os.path.exists(_lib.LIB_PATH), "run `python -m gcnn_keras_b200.build` first"
    handle = _lib.lib()
    assert handle.kgcnn_version() >= 100
    assert handle.kgcnn_error_string(0) == b"ok"
    assert handle.kgcnn_error_string(-3) == b"CUDA launch error"


def test_every_header_symbol_is_exported_and_bound():
    declared = _lib.header_symbols()
    assert len(declared) >= 15
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), "header declares %s but the library does not export it" % name
    assert set(declared) == set(_lib.SIGNATURES), \
        "ctypes table and header disagree: %s" % (set(declared) ^ set(_lib.SIGNATURES))


def test_argument_validation_without_gpu():
    h = _lib.lib()
    # invalid arguments are rejected before any launch, so these are safe on a CPU-only machine
    assert h.kgcnn_gather_rows(None, 4, 0, None, 1, 4, None, 4, None) == -1          # F <= 0
    assert h.kgcnn_gather_rows(None, 4, 8, None, 1, 0, None, 8, None) == 0           # E == 0 is a no-op
    assert h.kgcnn_gather_mul_segment_reduce(None, 8, None, 3, None, None, None, 1, 0, 4, 9, 0, None, None, None, 0, 0, None) == -1
    assert h.kgcnn_gather_mul_segment_reduce(None, 8, None, 0, None, None, None, 1, 0, 4, 9, 7, None, None, None, 0, 0, None) == -1
    assert h.kgcnn_gather_mul_segment_reduce(None, 8, None, 0, None, None, None, 1, 0, 0, 0, 0, None, None, None, 0, 0, None) == 0
    assert h.kgcnn_plan_workspace_bytes(1000, 100) > 3 * 4000


def test_no_cpu_fallback():
    from gcnn_keras_b200 import functional as F
    x = torch.zeros(4, 8)
    ids = torch.zeros(3, dtype=torch.int32)
    with pytest.raises(_lib.KgcnnError):
        F.raw_gather_rows(x, ids)


def test_product_does_not_import_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "gcnn_keras_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, \
                    "%s imports the oracle" % os.path.join(dirpath, f)


def test_segment_name_dispatch_errors():
    from gcnn_keras_b200 import functional as F
    assert F.op_code("segment_sum") == F.op_code("sum") == F.op_code("reduce_sum") == 0
    assert F.op_code("mean") == 1 and F.op_code("segment_max") == 2 and F.op_code("reduce_min") == 3
    with pytest.raises(TypeError):
        F.op_code("softmax")


def test_registered_layer_names_equal_the_reference_names():
    """Every layer registered here carries the Keras registration name the reference gives the class of the same
    name (tests/golden/registered_names.json, extracted from the reference's decorators by make_golden_names.py),
    so a serialized reference model config resolves to the right class."""
    import importlib
    import json
    import os
    import pkgutil

    import gcnn_keras_b200
    from gcnn_keras_b200.layers.base import _REGISTRY, get_registered
    for m in pkgutil.walk_packages(gcnn_keras_b200.__path__, "gcnn_keras_b200."):
        if m.name.endswith((".build", ".libkgcnn_b200", ".libkgcnn_b200_precise")):
            continue
        importlib.import_module(m.name)
    with open(os.path.join(os.path.dirname(__file__), "golden", "registered_names.json")) as fh:
        ref = json.load(fh)
    assert len(_REGISTRY) >= 40
    for key, cls in _REGISTRY.items():
        assert ref.get(cls.__name__) == key, (key, cls.__name__, ref.get(cls.__name__))
    assert get_registered("kgcnn>BesselBasisLayer").__name__ == "BesselBasisLayer"   # kgcnn/layers/geom.py:716


def test_peer_all_reduce_is_not_chosen_without_an_nccl_group():
    """distributed.PeerAllReduce.create: no process group (or gloo, or one rank) -> None, the caller keeps its
    collective; the kernel entry itself rejects bad arguments without touching a device."""
    import ctypes
    from gcnn_keras_b200 import _lib
    from gcnn_keras_b200.distributed import PeerAllReduce
    assert PeerAllReduce.create(1024, "cpu") is None
    assert PeerAllReduce.MAX_WORLD_AUTO == 4
    lib = _lib.lib()
    assert lib.kgcnn_allreduce_peer(None, 2, 16, 1.0, None, None) != 0
    arr = (ctypes.c_void_p * 1)(None)
    assert lib.kgcnn_allreduce_peer(arr, 1, 0, 1.0, ctypes.c_void_p(16), None) == 0      # n == 0: nothing to do
    assert lib.kgcnn_allreduce_peer(arr, 17, 16, 1.0, ctypes.c_void_p(16), None) != 0    # more ranks than slots


def test_static_weight_flag_needs_the_owner_to_opt_in():
    """cfconv stages its weights before the launch-dependency wait only for tensors inside a buffer whose owner has
    opted in (functional.static_weights); a copy made on the way to the kernel never qualifies."""
    from gcnn_keras_b200 import functional as F
    flat = torch.zeros(256)
    w, b = flat[:128].view(8, 16), flat[128:144]
    assert F.static_weight_flag((w, b, None)) == 0                      # default: wait first
    with F.static_weights(flat):
        assert F.static_weight_flag((w, b, None)) == 1
        assert F.static_weight_flag((w, b.clone())) == 0                # a fresh copy lies outside the buffer
        assert F.static_weight_flag((w, torch.zeros(4))) == 0
        with F.static_weights(None):                                    # no-op, the outer owner stays
            assert F.static_weight_flag((w,)) == 1
    assert F.static_weight_flag((w, b)) == 0
    assert F.static_weight_flag((torch.nn.Parameter(torch.zeros(3)),)) == 0   # being a leaf variable is not enough
