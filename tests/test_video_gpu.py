"""File-to-file entry points on the GPU (app.py:388-519 process_video, basic_upscaler_fallback.py:67-169
enhance_video, realesrgan_handler.py:214-273 restore_video) through the threaded reader/writer pipeline.
The clips are written here with OpenCV's own mp4v encoder; the codec is lossy, so frame content is compared with a
loose tolerance against the same decoded frames pushed through the frame API, while counts, sizes, fps and the
stats dictionaries are exact."""
import os

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from video_enhancer_proto_b200 import B200VideoEnhancer, synth

pytestmark = pytest.mark.gpu


def write_clip(path, n, w, h, fps=24.0):
    wr = cv2.VideoWriter(str(path), cv2.VideoWriter_fourcc(*"mp4v"), fps, (w, h))
    if not wr.isOpened():
        pytest.skip("OpenCV cannot write mp4v here")
    rng = np.random.default_rng(0)
    for i in range(n):
        wr.write(synth.demo_frame(i, size=(w, h), rng=rng, bgr=True))
    wr.release()


def read_clip(path):
    cap = cv2.VideoCapture(str(path))
    assert cap.isOpened()
    fps = cap.get(cv2.CAP_PROP_FPS)
    frames = []
    while True:
        ok, f = cap.read()
        if not ok:
            break
        frames.append(f)
    cap.release()
    return frames, fps


@pytest.mark.parametrize("model_name", ["AdvancedUpscaler", "bicubic"])
def test_enhance_video_file_to_file(tmp_path, model_name):
    src, dst = tmp_path / "in.mp4", tmp_path / "sub" / "out.mp4"
    n, w, h = 19, 96, 64                                   # 19 frames: two full batches of 8 and a ragged one
    write_clip(src, n, w, h)
    enh = B200VideoEnhancer(device="cuda:0", model_name=model_name, seed=0)
    try:
        stats = enh.enhance_video(str(src), str(dst))
        assert stats["frames_processed"] == n and stats["output_frames"] == n and stats["input_frames"] == n
        assert stats["input_resolution"] == f"{w}x{h}" and stats["output_resolution"] == f"{2 * w}x{2 * h}"
        assert stats["scale"] == 2 and stats["frames_failed"] == 0
        assert os.path.exists(dst)                          # parent directory is created (basic_upscaler_fallback.py:98)
        got, fps = read_clip(dst)
        ins, _ = read_clip(src)
        assert len(got) == n and got[0].shape == (2 * h, 2 * w, 3) and abs(fps - 24.0) < 0.5
        # content: decoded input frames through the frame API, compared with the decoded output (lossy codec)
        mode_flags = (0, 1) if model_name == "bicubic" else (1, 0b11)
        want = enh._run_host(mode_flags[0], ins, mode_flags[1]).copy()
        for i in (0, 7, 8, n - 1):                          # batch boundaries and the ragged tail, in order
            err = np.abs(got[i].astype(np.float32) - want[i].astype(np.float32)).mean()
            assert err < 8.0, (i, err)
        # order: frame i of the output is closer to its own expected frame than to its neighbours'
        d_own = np.abs(got[9].astype(np.float32) - want[9].astype(np.float32)).mean()
        d_far = np.abs(got[9].astype(np.float32) - want[0].astype(np.float32)).mean()
        assert d_own <= d_far
    finally:
        enh.close()


def test_process_video_and_restore_video(tmp_path):
    src = tmp_path / "in.mp4"
    write_clip(src, 10, 64, 48, fps=60.0)
    enh = B200VideoEnhancer(device="cuda:0", seed=0)
    try:
        ok, msg = enh.process_video(str(src), str(tmp_path / "o1.mp4"), target_fps=30)
        assert ok and msg.startswith("✅ Enhanced 10 frames")
        _, fps = read_clip(tmp_path / "o1.mp4")
        assert abs(fps - 30.0) < 0.5                        # writer fps = min(target_fps, fps) (app.py:447)
        assert enh.processing_stats["total_processed"] == 1 and len(enh.job_history) == 1
        s = enh.restore_video(str(src), str(tmp_path / "o2.mp4"))
        assert s["processing_mode"] == "frame_batch" and s["scale_factor"] == 2 and s["output_frames"] == 10
        ok, msg = enh.process_video(str(tmp_path / "missing.mp4"), str(tmp_path / "o3.mp4"))
        assert not ok and msg.startswith("❌")
        with pytest.raises(ValueError, match="Could not open video"):
            enh.enhance_video(str(tmp_path / "missing.mp4"), str(tmp_path / "o4.mp4"))
    finally:
        enh.close()
