#!/usr/bin/env python
"""Does OpenCV's mp4v writer get faster with encoder threads (OPENCV_FFMPEG_WRITER_OPTIONS)?  Encode-only, 2160p."""
import os, sys, time, subprocess
if len(sys.argv) > 1:
    opt = sys.argv[1]
    if opt != "default":
        os.environ["OPENCV_FFMPEG_WRITER_OPTIONS"] = opt
    import cv2, numpy as np
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from video_enhancer_proto_b200 import synth
    rng = np.random.default_rng(0)
    base = [np.ascontiguousarray(np.repeat(np.repeat(synth.demo_frame(i, size=(1920, 1080), rng=rng, bgr=True), 2, 0), 2, 1)) for i in range(6)]
    wr = cv2.VideoWriter("/tmp/enc_test.mp4", cv2.VideoWriter_fourcc(*"mp4v"), 30.0, (3840, 2160))
    assert wr.isOpened()
    t0 = time.perf_counter()
    for i in range(90): wr.write(base[i % 6])
    wr.release()
    dt = time.perf_counter() - t0
    cap = cv2.VideoCapture("/tmp/enc_test.mp4"); n = 0
    while cap.read()[0]: n += 1
    print("%-40s %.1f fps, file %.1f MB, %d frames read back" % (opt, 90 / dt, os.path.getsize("/tmp/enc_test.mp4") / 1e6, n), flush=True)
else:
    print("cpu count", os.cpu_count())
    for opt in ("default", "threads;4", "threads;8", "threads;16", "threads;16|thread_type;slice", "threads;0"):
        subprocess.run([sys.executable, __file__, opt])
