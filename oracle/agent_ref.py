"""CPU restatement of the reference agent's dispatch for the upscaler handlers.

TEST INFRASTRUCTURE - see ``oracle/__init__.py``.  Follows ``/root/reference/agents/enhancer/video_enhancer_sota.py``:
``_select_model`` (:275-314), ``_execute_sota_enhancement`` (:316-366) and ``_handle_video_enhancement`` (:155-210),
plus the one branch ``INTEGRATION.md`` §4 adds for the B200 handler.  ``tests/test_agent_registration.py`` pins it to the
reference's own class (AST-extracted from that file and patched with the same branch) wherever ``/root/reference`` exists;
the ``-m gpu`` test then drives the real ``B200VideoEnhancer`` through this dispatcher on the GPU box, where the reference
tree is absent.
"""
from __future__ import annotations

import enum
import os
import time
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, Optional

B200_MODEL_KEY = "b200_upscaler"          # the key INTEGRATION.md §4 registers in model_strategy


class TaskStatus(enum.Enum):              # agents/core/base_agent.py: the two states this path produces
    COMPLETED = "completed"
    FAILED = "failed"


@dataclass
class ProcessingResult:                   # agents/core/base_agent.py:105-113, same field names and defaults
    task_id: str
    status: TaskStatus
    output_data: Dict[str, Any] = field(default_factory=dict)
    error_message: Optional[str] = None
    processing_time: float = 0.0
    quality_metrics: Dict[str, float] = field(default_factory=dict)
    metadata: Dict[str, Any] = field(default_factory=dict)


class AgentRef:
    """The slice of ``VideoEnhancementAgent`` that leads from a task to a handler call."""

    def __init__(self, handlers: Dict[str, Any]):
        # video_enhancer_sota.py:81-86 constructs one handler object per model; here they are injected
        self.vsrm_handler = handlers.get("vsrm")
        self.seedvr2_handler = handlers.get("seedvr2")
        self.ditvr_handler = handlers.get("ditvr")
        self.fast_mamba_handler = handlers.get("fast_mamba_vsr")
        self.realesrgan_handler = handlers.get("realesrgan")
        self.rife_handler = handlers.get("rife")
        self.b200_handler = handlers.get(B200_MODEL_KEY)
        self.enhancement_stats = {                                            # :109-119 (+ the new key)
            "total_enhancements": 0, "successful_enhancements": 0, "failed_enhancements": 0,
            "model_usage": {"vsrm": 0, "seedvr2": 0, "ditvr": 0, "fast_mamba_vsr": 0, "realesrgan": 0, "rife": 0,
                            B200_MODEL_KEY: 0},
        }

    def _select_model(self, analysis_result, user_preferences) -> str:        # :275-314
        if user_preferences.get("force_model"):
            return user_preferences["force_model"]
        if user_preferences.get("fast_mode", False):
            return "fast_mamba_vsr"
        quality_score = analysis_result.get("quality_score", 0.5)
        motion_intensity = analysis_result.get("motion_intensity", 0.5)
        degradation_type = analysis_result.get("degradation_type", "unknown")
        task_type = analysis_result.get("task_type", "enhancement")
        if task_type == "interpolation" or user_preferences.get("interpolation", False):
            return "rife"
        if quality_score < 0.4:
            return "seedvr2"
        if task_type == "super_resolution" or user_preferences.get("super_resolution", False):
            return "realesrgan"
        if degradation_type == "unknown" or degradation_type not in ["blur", "noise", "compression"]:
            return "ditvr"
        if motion_intensity > 0.7:
            return "vsrm"
        return "realesrgan"

    def _execute_sota_enhancement(self, video_path, output_path, model_name, analysis_result):   # :316-366
        if model_name == "vsrm":
            return self.vsrm_handler.enhance_video(input_path=video_path, output_path=output_path, window=7, fp16=True)
        elif model_name == "seedvr2":
            return self.seedvr2_handler.restore_video(input_path=video_path, output_path=output_path,
                                                      quality_threshold=0.6)
        elif model_name == "ditvr":
            return self.ditvr_handler.restore_video(input_path=video_path, output_path=output_path,
                                                    degradation_type=analysis_result.get("degradation_type", "unknown"),
                                                    auto_adapt=True)
        elif model_name == "fast_mamba_vsr":
            return self.fast_mamba_handler.enhance_video(input_path=video_path, output_path=output_path, chunk_size=16,
                                                         fp16=True)
        elif model_name == "realesrgan":
            return self.realesrgan_handler.restore_video(input_path=video_path, output_path=output_path)
        elif model_name == "rife":
            return self.rife_handler.interpolate_video(input_path=video_path, output_path=output_path,
                                                       interpolation_factor=2)
        elif model_name == B200_MODEL_KEY:                                    # the branch INTEGRATION.md §4 adds
            return self.b200_handler.enhance_video(input_path=video_path, output_path=output_path)
        else:
            raise ValueError(f"Unknown model: {model_name}")

    def _generate_output_path(self, input_path: str) -> str:                  # :368-371
        p = Path(input_path)
        return str(p.parent / f"{p.stem}_sota_enhanced{p.suffix}")

    def handle_video_enhancement(self, task) -> ProcessingResult:             # :155-210 (the coroutine, run synchronously)
        start_time = time.time()
        input_data = task.input_data
        video_path = input_data.get("video_path", "")
        output_path = input_data.get("output_path", "")
        analysis_result = input_data.get("analysis_result", {})
        user_preferences = input_data.get("user_preferences", {})
        if not video_path or not os.path.exists(video_path):                  # :165-171
            return ProcessingResult(task_id=task.task_id, status=TaskStatus.FAILED,
                                    error_message=f"Video file not found: {video_path}",
                                    processing_time=time.time() - start_time)
        if not output_path:
            output_path = self._generate_output_path(video_path)
        try:
            self.enhancement_stats["total_enhancements"] += 1
            selected_model = self._select_model(analysis_result, user_preferences)
            enhancement_result = self._execute_sota_enhancement(video_path, output_path, selected_model, analysis_result)
            processing_time = time.time() - start_time
            self.enhancement_stats["successful_enhancements"] += 1
            self.enhancement_stats["model_usage"][selected_model] += 1
            return ProcessingResult(                                          # :194-204
                task_id=task.task_id, status=TaskStatus.COMPLETED, output_data=enhancement_result,
                processing_time=processing_time,
                metadata={"model_used": selected_model, "enhancement_type": "2025_sota",
                          "frames_processed": enhancement_result.get("output_frames", 0)})
        except Exception:
            self.enhancement_stats["failed_enhancements"] += 1                # :206-209 re-raises
            raise
