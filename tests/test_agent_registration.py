"""the handler registered in the reference's agent, executed rather than described.

* CPU (where ``/root/reference`` exists): the reference's ``VideoEnhancementAgent`` is AST-extracted from
  ``agents/enhancer/video_enhancer_sota.py``, the edit of ``INTEGRATION.md`` §4 is applied as an AST patch (one ``elif``
  in ``_execute_sota_enhancement``, one attribute, one statistics key), and ITS ``_handle_video_enhancement`` coroutine
  is run with recording stub handlers.  The restatement ``oracle/agent_ref.py`` must make the same decisions, the same
  handler calls and the same ``ProcessingResult`` for every case - that pins it.
* GPU: the restated dispatcher drives the real ``B200VideoEnhancer`` on a clip and the ``ProcessingResult`` metadata the
  agent derives from the stats dictionary (``frames_processed`` <- ``output_frames``, :202) is checked.
"""
import ast
import asyncio
import os
import types

import numpy as np
import pytest

from oracle import agent_ref as A

REF = "/root/reference"
AGENT_PY = os.path.join(REF, "agents", "enhancer", "video_enhancer_sota.py")

PATCH_ELIF = """
if model_name == 'b200_upscaler':
    return self.b200_handler.enhance_video(input_path=video_path, output_path=output_path)
"""


class Recorder:
    """Stand-in handler: records the call, returns a stats dictionary."""

    def __init__(self, name, log, stats):
        self.name, self.log, self.stats = name, log, stats

    def _call(self, method, **kw):
        self.log.append((self.name, method, tuple(sorted(kw.items()))))
        return dict(self.stats)

    def enhance_video(self, **kw):
        return self._call("enhance_video", **kw)

    def restore_video(self, **kw):
        return self._call("restore_video", **kw)

    def interpolate_video(self, **kw):
        return self._call("interpolate_video", **kw)


def b200_stats(n=19, w=96, h=64):
    """The dictionary B200VideoEnhancer.enhance_video returns, built by the class's own code (no GPU needed)."""
    from video_enhancer_proto_b200 import B200VideoEnhancer
    s = B200VideoEnhancer.video_stats(n, w, h, 2, "AdvancedUpscaler", 24.0, {"fps": 24.0})
    s.pop("_size")
    return s


def reference_agent(log, stats):
    """The reference class, extracted and patched; __init__ is bypassed (it constructs six heavyweight handlers)."""
    import sys
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from agents.core import (BaseVideoProcessingAgent, AgentCapabilities, TaskSpecification, ProcessingResult,
                             TaskStatus, Priority)
    tree = ast.parse(open(AGENT_PY).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "VideoEnhancementAgent")
    fn = next(n for n in cls.body if isinstance(n, ast.AsyncFunctionDef) and n.name == "_execute_sota_enhancement")
    chain = next(n for n in fn.body if isinstance(n, ast.If))
    while chain.orelse and isinstance(chain.orelse[0], ast.If):            # walk to the last elif
        chain = chain.orelse[0]
    new_if = ast.parse(PATCH_ELIF).body[0]
    new_if.orelse = chain.orelse                                           # keeps the final `else: raise ValueError`
    chain.orelse = [new_if]
    mod = ast.Module(body=[cls], type_ignores=[])
    ast.fix_missing_locations(mod)
    import logging, time as _time
    from pathlib import Path
    from typing import Any, Dict, List, Optional, Tuple
    ns = dict(BaseVideoProcessingAgent=BaseVideoProcessingAgent, AgentCapabilities=AgentCapabilities,
              TaskSpecification=TaskSpecification, ProcessingResult=ProcessingResult, TaskStatus=TaskStatus,
              Priority=Priority, logger=logging.getLogger("ref_agent"), time=_time, os=os, Path=Path, Dict=Dict, Any=Any,
              List=List, Optional=Optional, Tuple=Tuple, asyncio=asyncio)
    exec(compile(mod, AGENT_PY, "exec"), ns)
    agent = ns["VideoEnhancementAgent"].__new__(ns["VideoEnhancementAgent"])
    for attr, key in (("vsrm_handler", "vsrm"), ("seedvr2_handler", "seedvr2"), ("ditvr_handler", "ditvr"),
                      ("fast_mamba_handler", "fast_mamba_vsr"), ("realesrgan_handler", "realesrgan"),
                      ("rife_handler", "rife"), ("b200_handler", A.B200_MODEL_KEY)):
        setattr(agent, attr, Recorder(key, log, stats))
    agent.enhancement_stats = {"total_enhancements": 0, "successful_enhancements": 0, "failed_enhancements": 0,
                               "model_usage": {k: 0 for k in ("vsrm", "seedvr2", "ditvr", "fast_mamba_vsr",
                                                              "realesrgan", "rife", A.B200_MODEL_KEY)}}
    return agent


CASES = [
    ({}, {"force_model": A.B200_MODEL_KEY}),
    ({}, {"fast_mode": True}),
    ({"task_type": "interpolation"}, {}),
    ({"quality_score": 0.2}, {}),
    ({"task_type": "super_resolution", "degradation_type": "blur"}, {}),
    ({"degradation_type": "blur", "motion_intensity": 0.9}, {}),
    ({"degradation_type": "noise"}, {}),
    ({}, {}),
]


@pytest.mark.skipif(not os.path.exists(AGENT_PY), reason="reference tree not present (GPU box)")
def test_restated_dispatch_equals_the_patched_reference_agent(tmp_path):
    clip = tmp_path / "clip.mp4"
    clip.write_bytes(b"x")                                                 # the agent only checks that the file exists (:165)
    stats = b200_stats()
    for analysis, prefs in CASES:
        task = types.SimpleNamespace(task_id="t1", task_type="video_enhancement",
                                     input_data={"video_path": str(clip), "output_path": "", "analysis_result": analysis,
                                                 "user_preferences": prefs})
        log_ref, log_mine = [], []
        ref = reference_agent(log_ref, stats)
        mine = A.AgentRef({k: Recorder(k, log_mine, stats) for k in
                           ("vsrm", "seedvr2", "ditvr", "fast_mamba_vsr", "realesrgan", "rife", A.B200_MODEL_KEY)})
        r_ref = asyncio.run(ref._handle_video_enhancement(task))
        r_mine = mine.handle_video_enhancement(task)
        assert log_ref == log_mine and len(log_ref) == 1, (analysis, prefs)
        assert r_ref.status.name == r_mine.status.name == "COMPLETED"
        assert r_ref.metadata == r_mine.metadata and r_ref.output_data == r_mine.output_data
        assert r_ref.metadata["frames_processed"] == stats["output_frames"] == 19        # :202 reads `output_frames`
        assert ref.enhancement_stats == mine.enhancement_stats
    # the B200 branch really is the one taken for its key, with the reference's keyword names
    assert log_ref[0][0] != A.B200_MODEL_KEY
    log = []
    ref = reference_agent(log, stats)
    task = types.SimpleNamespace(task_id="t2", task_type="video_enhancement",
                                 input_data={"video_path": str(clip), "output_path": "/o.mp4",
                                             "user_preferences": {"force_model": A.B200_MODEL_KEY}})
    r = asyncio.run(ref._handle_video_enhancement(task))
    assert log == [(A.B200_MODEL_KEY, "enhance_video", (("input_path", str(clip)), ("output_path", "/o.mp4")))]
    assert r.metadata["model_used"] == A.B200_MODEL_KEY
    # missing file and unknown model behave like the reference
    bad = types.SimpleNamespace(task_id="t3", task_type="video_enhancement", input_data={"video_path": "/nope.mp4"})
    assert asyncio.run(ref._handle_video_enhancement(bad)).status.name == "FAILED"
    assert A.AgentRef({}).handle_video_enhancement(bad).status.name == "FAILED"
    unk = types.SimpleNamespace(task_id="t4", task_type="video_enhancement",
                                input_data={"video_path": str(clip), "user_preferences": {"force_model": "nope"}})
    with pytest.raises(ValueError, match="Unknown model"):
        asyncio.run(ref._handle_video_enhancement(unk))
    with pytest.raises(ValueError, match="Unknown model"):
        A.AgentRef({}).handle_video_enhancement(unk)


def test_stats_dictionary_has_the_keys_the_reference_reads():
    s = b200_stats()
    for k in ("frames_processed", "output_frames", "input_frames", "scale", "model", "input_resolution",
              "output_resolution"):          # video_enhancer_sota.py:202, performance_monitor.py:497-505, basic_upscaler_fallback.py:154-162
        assert k in s
    assert s["input_resolution"] == "96x64" and s["output_resolution"] == "192x128"


@pytest.mark.gpu
def test_agent_dispatch_runs_the_b200_handler(tmp_path):
    cv2 = pytest.importorskip("cv2")
    from video_enhancer_proto_b200 import B200VideoEnhancer, synth
    src = tmp_path / "in.mp4"
    wr = cv2.VideoWriter(str(src), cv2.VideoWriter_fourcc(*"mp4v"), 24.0, (96, 64))
    if not wr.isOpened():
        pytest.skip("OpenCV cannot write mp4v here")
    rng = np.random.default_rng(0)
    for i in range(11):
        wr.write(synth.demo_frame(i, size=(96, 64), rng=rng, bgr=True))
    wr.release()
    handler = B200VideoEnhancer(device="cuda:0", seed=0)                  # constructed with device= only, like :81-86
    try:
        agent = A.AgentRef({A.B200_MODEL_KEY: handler})
        task = types.SimpleNamespace(task_id="job-7", task_type="video_enhancement",
                                     input_data={"video_path": str(src), "user_preferences": {"force_model": A.B200_MODEL_KEY}})
        r = agent.handle_video_enhancement(task)
        assert r.status is A.TaskStatus.COMPLETED and r.task_id == "job-7"
        assert r.metadata == {"model_used": A.B200_MODEL_KEY, "enhancement_type": "2025_sota", "frames_processed": 11}
        assert r.output_data["output_frames"] == 11 and r.output_data["output_resolution"] == "192x128"
        out = tmp_path / "in_sota_enhanced.mp4"                           # _generate_output_path (:368-371)
        assert out.exists()
        cap = cv2.VideoCapture(str(out))
        assert int(cap.get(cv2.CAP_PROP_FRAME_WIDTH)) == 192 and int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT)) == 128
        cap.release()
        assert agent.enhancement_stats["model_usage"][A.B200_MODEL_KEY] == 1
        info = handler.get_model_info()                                   # get_processing_stats (:377-384) reads this
        assert info["scale"] == 2 and "name" in info and "description" in info
    finally:
        handler.close()


@pytest.mark.gpu
def test_agent_realesrgan_branch_runs_the_b200_handler(tmp_path):
    """The agent's own 'realesrgan' branch (video_enhancer_sota.py:353-356) served by the GPU twin of RealESRGANHandler:
    `restore_video(input_path=, output_path=)` -> the statistics dictionary of realesrgan_handler.py:265-273."""
    cv2 = pytest.importorskip("cv2")
    from video_enhancer_proto_b200 import B200RealESRGAN, synth
    src = tmp_path / "in.mp4"
    wr = cv2.VideoWriter(str(src), cv2.VideoWriter_fourcc(*"mp4v"), 24.0, (64, 48))
    if not wr.isOpened():
        pytest.skip("OpenCV cannot write mp4v here")
    rng = np.random.default_rng(0)
    for i in range(5):
        wr.write(synth.demo_frame(i, size=(64, 48), rng=rng, bgr=True))
    wr.release()
    handler = B200RealESRGAN(device="cuda:0", scale=4, seed=0, num_block=1)
    try:
        agent = A.AgentRef({"realesrgan": handler})
        task = types.SimpleNamespace(task_id="job-8", task_type="video_enhancement",
                                     input_data={"video_path": str(src), "analysis_result": {"task_type": "super_resolution"}})
        r = agent.handle_video_enhancement(task)                        # _select_model picks 'realesrgan' (:297-299)
        assert r.status is A.TaskStatus.COMPLETED
        assert r.metadata == {"model_used": "realesrgan", "enhancement_type": "2025_sota", "frames_processed": 5}
        assert r.output_data["output_resolution"] == (192, 256) and r.output_data["scale_factor"] == 4
        assert (tmp_path / "in_sota_enhanced.mp4").exists()
    finally:
        handler.close()
