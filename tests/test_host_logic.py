"""CPU: host-side logic — weights procedure, frame sharding (incl. a world_size-2 gloo run), interface shape."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

from video_enhancer_proto_b200 import shard, weights as W

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_init_state_is_the_reference_procedure():
    from oracle import upscaler_ref as O
    for seed in (0, 3):
        a, b = W.init_state(seed), O.export_state(O.make_model(seed))
        assert list(a) == list(b) == list(W.STATE_DICT_KEYS)
        for k in a:
            assert np.array_equal(a[k], b[k]), k


def test_normalize_state_rejects_bad_shapes():
    st = dict(W.init_state(0))
    W.normalize_state(st)
    bad = dict(st)
    bad["conv3.weight"] = bad["conv3.weight"][:, :64]
    with pytest.raises(ValueError):
        W.normalize_state(bad)
    del bad["conv3.weight"]
    with pytest.raises(KeyError):
        W.normalize_state(bad)
    assert W.infer_plan(st) == ((128, 128, 256, 128), 64, 0)


def test_deepened_plan_state_matches_oracle_order_and_values():
    """configs[4]: extra conv3x<k>/bn3x<k> blocks are appended after the reference's 30 tensors, and the init
    procedure draws them in module order exactly like the oracle's model."""
    from oracle import upscaler_ref as O
    a = W.init_state(3, widths=(128, 128, 256, 128), n_extra=2)
    b = O.export_state(O.make_model(3, widths=(128, 128, 256, 128), extra=2))
    assert list(a) == list(b) and len(a) == 30 + 6 * 2
    assert list(a)[:30] == list(W.STATE_DICT_KEYS) and list(a)[30:32] == ["conv3x1.weight", "conv3x1.bias"]
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    assert W.infer_plan(a) == ((128, 128, 256, 128), 64, 2)
    assert a["conv3x2.weight"].shape == (256, 256, 3, 3)


@pytest.mark.parametrize("world,n", [(1, 5), (2, 5), (4, 3000), (8, 3000), (8, 3), (3, 0)])
def test_frame_ranges_partition(world, n):
    rs = shard.all_ranges(world, n)
    assert rs[0][0] == 0 and rs[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
    sizes = [hi - lo for lo, hi in rs]
    assert max(sizes) - min(sizes) <= 1
    for f in range(0, n, max(1, n // 7)):
        r = shard.owner_of(f, world, n)
        assert rs[r][0] <= f < rs[r][1]
    merged = shard.merge_ordered([list(range(lo, hi)) for lo, hi in rs], world, n)
    assert merged == list(range(n))
    if n:
        with pytest.raises(ValueError):
            shard.merge_ordered([[] for _ in rs], world, n)


def _worker(rank, world, port, n_frames, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard.frame_range(rank, world, n_frames)
    # each rank "processes" its contiguous range; the per-frame result is a tag derived from the index
    local = [(i, (i * 2654435761) & 0xFFFFFFFF) for i in range(lo, hi)]
    merged = shard.gather_to_rank0(local, n_frames)
    t = torch.tensor([float(hi - lo)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)              # the bench's max-over-ranks reduction
    if rank == 0:
        q.put((merged, t.item()))
    dist.barrier()
    dist.destroy_process_group()


def test_world2_gloo_sharding_and_host_reorder():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    n = 11
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    [p.start() for p in procs]
    merged, mx = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert [m[0] for m in merged] == list(range(n))
    assert all(m[1] == (m[0] * 2654435761) & 0xFFFFFFFF for m in merged)
    assert mx == 6.0
    assert all(p.exitcode == 0 for p in procs)


def test_interface_surface_matches_reference_callers():
    """Names/signatures the reference's callers use."""
    import inspect
    from video_enhancer_proto_b200 import B200VideoEnhancer, GPUVideoEnhancer
    assert GPUVideoEnhancer is B200VideoEnhancer
    sig = inspect.signature(B200VideoEnhancer.__init__)
    assert list(sig.parameters)[1:4] == ["device", "model_name", "scale"]       # basic_upscaler_fallback.py:45
    assert list(inspect.signature(B200VideoEnhancer.process_video).parameters)[1:] == ["input_path", "output_path", "target_fps"]
    assert inspect.signature(B200VideoEnhancer.process_video).parameters["target_fps"].default == 30   # app.py:388
    for m in ("enhance_video", "restore_video", "get_model_info", "load_models", "enhance_frame_gpu",
              "_enhance_frame_with_gpu", "_enhance_frame_cpu_fallback", "_process_frame_batch_gpu"):
        assert callable(getattr(B200VideoEnhancer, m))
    e = B200VideoEnhancer(device="cuda")
    info = e.get_model_info()
    assert info["scale"] == 2 and info["parameters"] == 750982
    assert e.models_loaded is False
    with pytest.raises(ValueError):
        B200VideoEnhancer(scale=4)


def test_synth_generators_are_deterministic():
    from video_enhancer_proto_b200 import synth
    a = synth.mixed_clip(6, 48, 64, seed=5)
    b = synth.mixed_clip(6, 48, 64, seed=5)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert all(f.shape == (48, 64, 3) and f.dtype == np.uint8 for f in a)
    t = synth.tag_frame(a[0], 0x01020304)
    assert synth.read_tag(t) == 0x01020304


def test_basic_upscaler_twin_constructor_semantics():
    """basic_upscaler_fallback.py:45-62: (device, model_name, scale) order, unknown names fall back to Lanczos; all five
    interpolation names of the reference's table (:51-57) map to a kernel.  No CUDA is touched before the first frame."""
    import inspect
    from video_enhancer_proto_b200.basic_upscaler import B200BasicUpscaler, SHARPEN_KERNEL
    assert list(inspect.signature(B200BasicUpscaler.__init__).parameters)[1:4] == ["device", "model_name", "scale"]
    assert B200BasicUpscaler(model_name="cubic").method == "CUBIC"
    assert B200BasicUpscaler(model_name="something-else").method == "LANCZOS"
    for name in ("LINEAR", "area", "NEAREST"):
        assert B200BasicUpscaler(model_name=name).method == name.upper()
    assert B200BasicUpscaler.CV_INTERPOLATION == {"NEAREST": 0, "LINEAR": 1, "CUBIC": 2, "AREA": 3, "LANCZOS": 4}
    with pytest.raises(ValueError):
        B200BasicUpscaler(scale=4)
    info = B200BasicUpscaler(model_name="LANCZOS").get_model_info()
    assert {"name", "description", "scale", "device", "interpolation_method", "capabilities"} <= set(info)
    # basic_upscaler_fallback.py:123-124 in float32: the whole kernel is scaled, so it sums to 0.1
    assert SHARPEN_KERNEL.dtype == np.float32 and abs(float(SHARPEN_KERNEL.sum()) - 0.1) < 1e-6
    assert float(SHARPEN_KERNEL[1, 1]) == float(np.float32(9) * np.float32(0.1))


def test_bench_reads_roofline_traffic_from_the_committed_ncu_summaries(tmp_path, monkeypatch):
    """`roofline.traffic` must follow the ncu capture named in `traffic_source`, not a literal."""
    sys.path.insert(0, ROOT)
    import bench
    b, src, n, per = bench.ncu_traffic(bench.PROFILE_CNN_CSV, r"conv3x3_tc_kernel")
    assert src and os.path.exists(os.path.join(ROOT, src)) and n == 3 and len(per) == 3
    assert 5e9 < b < 3e10                                   # GB-scale per launch at batch 8
    b2, src2, n2, _ = bench.ncu_traffic(bench.PROFILE_BIC_CSV, r"bicubic2x_strip")
    assert src2 and n2 >= 1 and 9.33e9 <= b2 < 1.1 * 9.33e9   # at least the algorithmic bytes of 300 frames
    # units are honoured and a newer file wins
    p = tmp_path / "profiles"
    p.mkdir()
    (p / "x.csv").write_text('"Kernel Name","dram__bytes_read.sum","dram__bytes_write.sum"\n"","Mbyte","Gbyte"\n'
                             '"k1(int)","500","1.5"\n"k1(int)","700","1.5"\n"other","1","1"\n')
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    b3, src3, n3, per3 = bench.ncu_traffic(("profiles/missing.csv", "profiles/x.csv"), r"k1")
    assert src3 == "profiles/x.csv" and n3 == 2 and abs(b3 - 2.1e9) < 1 and list(per3) == ["k1"]
    assert bench.ncu_traffic(("profiles/missing.csv",), r"k1") == (None, None, 0, {})
