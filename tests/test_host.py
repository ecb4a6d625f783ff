"""CPU-side checks: the drop-in module contract, the C-ABI library's exports, error behaviour
and the multi-GPU sharding helpers (gloo, world_size 2). No compute kernels run here."""
import contextlib
import ctypes
import io
import json
import os
import pickle
import re
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _model():
    from speechseparation_b200 import BSRNN
    with contextlib.redirect_stdout(io.StringIO()):
        return BSRNN()


def test_library_exports_every_header_symbol():
    from speechseparation_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "bsrnn_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(bsb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    l = _lib.lib()
    for name in declared:
        assert hasattr(l, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert l.bsb_version() >= 100
    assert l.bsb_error_string(-3).decode().endswith("workspace too small")


def test_library_tables_match_reference_contract(golden_dir):
    from speechseparation_b200 import _lib, generate_bandsplits
    l = _lib.lib()
    c = json.load(open(os.path.join(golden_dir, "contract.json")))
    v, w = generate_bandsplits()
    assert v == c["bandsplits"]
    assert w == [3, 5, 9, 18, 36, 72, 144, 288, 576, 641, 257, 1]
    numel = [l.bsb_tensor_numel(i) for i in range(288)]
    ref = [int(torch.Size(s).numel()) for _, s in c["keys"]]
    assert numel == ref and l.bsb_tensor_numel(288) == -1
    assert l.bsb_packed_weights_bytes() >= 4 * c["n_params"]
    assert l.bsb_num_frames(441000) == 431 and l.bsb_num_frames(176400) == 173 and l.bsb_num_frames(1024) == -1
    assert l.bsb_workspace_bytes(2, 173) > 2 * 173 * 2050 * 4 * 4
    assert l.bsb_workspace_bytes(0, 5) == 0


def test_argument_errors_do_not_need_a_gpu():
    from speechseparation_b200 import _lib
    l = _lib.lib()
    assert l.bsb_forward(None, 2, 3, None, None, None, 0, None) == -1
    assert l.bsb_separate(None, 1, 4096, None, None, None, None, 0, None) == -1
    assert l.bsb_pack_weights(None, 288, None, None) == -1
    table = (ctypes.c_void_p * 10)()
    assert l.bsb_pack_weights(table, 10, ctypes.c_void_p(256), None) == -1


def test_state_dict_contract_and_default_init(golden_dir):
    c = json.load(open(os.path.join(golden_dir, "contract.json")))
    torch.manual_seed(0)
    m = _model()
    sd = m.state_dict()
    assert [[k, list(v.shape)] for k, v in sd.items()] == c["keys"]
    assert sum(v.numel() for v in sd.values()) == 7481062
    # same construction order => same RNG draws as `torch.manual_seed(0); reference.BSRNN()`
    if torch.__version__ == c["torch"]:
        for k, v in sd.items():
            assert abs(float(v.double().sum()) - c["seed0_sum"][k]) < 1e-9, k


def test_strict_load_and_pickle(oracle_sd, tmp_path):
    m = _model()
    m.load_state_dict(oracle_sd, strict=True)
    bad = dict(oracle_sd)
    bad.pop("bandFCs.11.0.trainable_constant")
    with pytest.raises(RuntimeError):
        _model().load_state_dict(bad, strict=True)
    p = tmp_path / "model-full.pth"
    torch.save(m, p)                                   # infer-streaming.py:49
    m2 = torch.load(p, weights_only=False)
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), m2.state_dict().values()))
    assert m2._packed is None and pickle.loads(pickle.dumps(m))._ws == {}
    assert len(list(m.parameters())) == 288 and m.eval() is m and m.train() is m


def test_no_cpu_fallback_and_shape_errors():
    from speechseparation_b200 import stft_pack, unpack_istft
    m = _model()
    with pytest.raises(RuntimeError, match="no CPU path"):
        m.forward(torch.zeros(2, 2050, 4))
    with pytest.raises(RuntimeError, match="no CPU path"):
        m.forward_recurrent(torch.zeros(2, 2050), torch.zeros(4, 2, 24, 64))
    with pytest.raises(RuntimeError):
        stft_pack(torch.zeros(2, 4096))
    with pytest.raises(RuntimeError):
        unpack_istft(torch.zeros(2, 2050, 4))
    with pytest.raises(RuntimeError):
        m.separate(torch.zeros(2, 4096))


def test_root_shim_imports_like_the_reference():
    sys.path.insert(0, ROOT)
    import importlib
    mod = importlib.import_module("bsrnn")
    assert mod.BSRNN.__module__ == "speechseparation_b200.bsrnn"
    assert mod.generate_bandsplits()[0][-2:] == [257, 0]


def test_shard_range_partitions_rows():
    from speechseparation_b200 import shard_range
    for n in (0, 1, 7, 64, 1024):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    from speechseparation_b200 import aggregate_throughput, shard_range
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    lo, hi = shard_range(1024, rank, world)              # cfg 4: 1024 session rows over G ranks
    units, secs = aggregate_throughput(float(hi - lo), 0.5 + rank, device="cpu")
    dist.barrier()
    q.put((rank, lo, hi, units, secs))
    dist.destroy_process_group()


def test_gloo_world2_sharding_and_aggregation():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29000 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert [(r[1], r[2]) for r in res] == [(0, 512), (512, 1024)]
    assert all(r[3] == 1024.0 and r[4] == 1.5 for r in res)   # sum of units, max of seconds


def test_plugin_header_symbols_and_ladspa_descriptor():
    """include/bsrnn_b200_plugin.h: every declared function is exported; the LADSPA descriptor carries the
    reference plugin's identity and port layout (speech-ladspa-onnx.cpp:293-337). No GPU needed."""
    from speechseparation_b200 import _lib, plugin
    hdr = open(os.path.join(ROOT, "include", "bsrnn_b200_plugin.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(bsb_plugin_[a-z0-9_]+|ladspa_descriptor)\s*\(", hdr))
    l = _lib.lib()
    for name in declared:
        assert hasattr(l, name), f"{name} declared in the plugin header but not exported"
    assert declared == set(_lib.PLUGIN_SIGNATURES), declared ^ set(_lib.PLUGIN_SIGNATURES)
    d = plugin.ladspa_descriptor(0)
    assert plugin.ladspa_descriptor(1) is None
    assert d.UniqueID == 0xf433b044 and d.Label == b"speech_separator" and d.PortCount == 5 and d.Properties == 0
    assert [d.PortDescriptors[i] for i in range(5)] == [0x1 | 0x4, 0x1 | 0x8, 0x1 | 0x8, 0x2 | 0x8, 0x2 | 0x8]
    assert [d.PortNames[i] for i in range(5)] == [b"Control", b"Input (Left)", b"Input (Right)", b"Output (Left)",
                                                  b"Output (Right)"]
    assert d.PortRangeHints[0].HintDescriptor == (0x240 | 0x1) and d.PortRangeHints[0].LowerBound == 0.0
    assert all(d.PortRangeHints[i].HintDescriptor == 0 for i in range(1, 5))
    assert not d.activate and not d.run_adding and not d.deactivate
    assert d.instantiate and d.connect_port and d.run and d.cleanup
    assert l.bsb_plugin_weight_floats() == 7481062


def test_plugin_argument_errors_and_weight_file(tmp_path, oracle_sd):
    from speechseparation_b200 import _lib, plugin
    l = _lib.lib()
    h = ctypes.c_void_p()
    assert l.bsb_plugin_create(ctypes.byref(h), None, 7481062, 0) == -1 and not h
    w = plugin.flat_weights(oracle_sd)
    assert w.size == 7481062 and w.dtype.name == "float32"
    assert l.bsb_plugin_create(ctypes.byref(h), w.ctypes.data, 10, 0) == -1 and not h
    assert l.bsb_plugin_run(None, 4) == -1 and l.bsb_plugin_hops(None) == -1
    short = tmp_path / "short.bin"
    w[:100].tofile(short)
    assert l.bsb_plugin_create_from_file(ctypes.byref(h), str(short).encode(), 0) == -2      # wrong size: shape error
    assert l.bsb_plugin_create_from_file(ctypes.byref(h), b"/nonexistent/w.bin", 0) == -1
    # flat image = state_dict order
    first = next(iter(oracle_sd.values())).reshape(-1).numpy()
    assert (w[:first.size] == first).all()
    with pytest.raises(RuntimeError):
        plugin.flat_weights({"a": torch.zeros(3)})


def test_export_operators_registered_with_shapes_and_no_cpu_kernel():
    """the two model entry points exist as torch.library operators (what a tracer records);
    shape propagation works without a device and there is no CPU kernel behind them."""
    import speechseparation_b200  # noqa: F401  (registers the operators)
    s = str(torch.ops.bsrnn_b200.forward_recurrent.default._schema)
    assert s == "bsrnn_b200::forward_recurrent(Tensor x, Tensor state, Tensor packed) -> (Tensor, Tensor)"
    assert str(torch.ops.bsrnn_b200.forward.default._schema) == "bsrnn_b200::forward(Tensor x, Tensor packed) -> Tensor"
    x, st, pk = (torch.empty(s, device="meta") for s in ((2, 2050), (4, 2, 24, 64), (16,)))
    y, ns = torch.ops.bsrnn_b200.forward_recurrent(x, st, pk)
    assert y.shape == (2, 2050) and ns.shape == (4, 2, 24, 64)
    assert torch.ops.bsrnn_b200.forward(torch.empty(3, 2050, 7, device="meta"), pk).shape == (3, 2050, 7)
    with pytest.raises(NotImplementedError):
        torch.ops.bsrnn_b200.forward_recurrent(torch.zeros(2, 2050), torch.zeros(4, 2, 24, 64), torch.zeros(16))
    with pytest.raises(NotImplementedError):
        torch.ops.bsrnn_b200.forward(torch.zeros(1, 2050, 2), torch.zeros(16))
