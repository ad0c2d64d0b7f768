"""CPU oracle for the per-frame 2x upscaler hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the shipped product path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and there only as the checker (or as the timed CPU baseline),
never as the thing that produces the product's output.

What it restates (all citations relative to ``/root/reference``):

* ``AdvancedUpscaler`` and its init loop          app.py:269-301, 303-311
* batch pre/post-processing                       app.py:533-551
* single-frame pre/post-processing                app.py:334-346
* bicubic CPU fallback                            app.py:364-376
* OpenCV bicubic twin                             models/enhancement/frame/basic_upscaler_fallback.py:142
* BasicUpscalerFallback per-frame body            models/enhancement/frame/basic_upscaler_fallback.py:113-128
* Real-ESRGAN network and handler frame path      models/enhancement/vsr/realesrgan_handler.py:23-100, 286-372
                                                  (oracle/realesrgan_ref.py, pinned by oracle/make_golden_rrdb.py)
* the agent's dispatch to the handlers            agents/enhancer/video_enhancer_sota.py:155-210, 275-371
                                                  (oracle/agent_ref.py, pinned by tests/test_agent_registration.py)

Parity pinning status: the reference's own tests hold NO golden vectors, known-answer
tests or tolerances for this path, so the restatement is pinned
against outputs of the reference code itself, executed in the build container
(``oracle/make_golden.py`` AST-extracts ``AdvancedUpscaler`` from app.py and runs the
verbatim pre/post lines); the resulting vectors live in ``tests/golden/``.
"""
