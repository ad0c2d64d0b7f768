"""Stream facts for the file-to-file calls, with the key set and fall-back order of the reference's
``VideoUtils.get_video_metadata`` (utils/video_utils.py:38-135): ``ffprobe`` when it is on PATH, otherwise the container's
own header through OpenCV; on any failure a zero-filled record with ``fps`` 30.  ``frame_count`` is an alias of
``total_frames`` (utils/video_utils.py:79,122).  Host-side only; nothing here touches CUDA."""
from __future__ import annotations

import json
import os
import shutil
import subprocess
from typing import Any, Dict

KEYS = ("duration", "width", "height", "fps", "total_frames", "frame_count", "codec", "format", "bitrate", "file_size",
        "pixel_format")


def _empty() -> Dict[str, Any]:
    return {"duration": 0, "width": 0, "height": 0, "fps": 30, "total_frames": 0, "frame_count": 0, "codec": "unknown",
            "format": "unknown", "bitrate": 0, "file_size": 0, "pixel_format": "unknown"}


def parse_rate(text: str) -> float:
    """'30000/1001' or '25' -> frames per second; 30.0 when unparsable (utils/video_utils.py:97-105)."""
    try:
        if "/" in text:
            num, den = text.split("/")
            return float(num) / float(den)
        return float(text)
    except (ValueError, ZeroDivisionError, AttributeError):
        return 30.0


def _from_ffprobe(path: str) -> Dict[str, Any] | None:
    exe = shutil.which("ffprobe")
    if not exe:
        return None
    res = subprocess.run([exe, "-v", "quiet", "-print_format", "json", "-show_format", "-show_streams", path],
                         capture_output=True, text=True)
    if res.returncode != 0:
        return None
    data = json.loads(res.stdout)
    stream = next((s for s in data.get("streams", []) if s.get("codec_type") == "video"), None)
    if stream is None:
        return None
    fmt = data.get("format", {})
    m = _empty()
    m.update(duration=float(fmt.get("duration", 0)), bitrate=int(fmt.get("bit_rate", 0)), width=int(stream.get("width", 0)),
             height=int(stream.get("height", 0)), fps=parse_rate(stream.get("r_frame_rate", "30/1")),
             codec=stream.get("codec_name", "unknown"), format=fmt.get("format_name", "unknown"),
             file_size=int(fmt.get("size", 0)), pixel_format=stream.get("pix_fmt", "unknown"))
    frames = int(m["fps"] * m["duration"]) if m["duration"] > 0 and m["fps"] > 0 else 0
    nb = stream.get("nb_frames")
    if nb and str(nb).isdigit():
        frames = int(nb)
    m["total_frames"] = m["frame_count"] = frames
    return m


def _from_opencv(path: str) -> Dict[str, Any]:
    import cv2

    cap = cv2.VideoCapture(path)
    try:
        n = int(cap.get(cv2.CAP_PROP_FRAME_COUNT))
        fps = cap.get(cv2.CAP_PROP_FPS)
        m = _empty()
        m.update(duration=n / fps if fps > 0 else 0, width=int(cap.get(cv2.CAP_PROP_FRAME_WIDTH)),
                 height=int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT)), fps=fps, total_frames=n, frame_count=n,
                 file_size=os.path.getsize(path) if os.path.exists(path) else 0)
        return m
    finally:
        cap.release()


def video_metadata(path: str) -> Dict[str, Any]:
    try:
        m = _from_ffprobe(str(path))
        return m if m is not None else _from_opencv(str(path))
    except Exception:
        try:
            return _from_opencv(str(path))
        except Exception:
            return _empty()
