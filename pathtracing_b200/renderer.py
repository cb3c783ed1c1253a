"""Host-side mirror of the reference's `Renderer` for the compute path only.

Same member names, argument meaning and error behaviour as the members of
RayTracing/Graphics/Renderer.cs that drive the tracing path, so the parity tests read like the
reference's call sequence:

    Renderer(IWindow)            Renderer.cs:59-64     -> Renderer(window)
    Init()                       Renderer.cs:66-84     -> CreateResources + CreateComputePipeline
    CreateResources()            Renderer.cs:105-196   -> allocates the W x H RGBA8 image (ptb_create)
    CreateComputePipeline()      Renderer.cs:293-403   -> binds the CSMain-equivalent kernel
    ComputeFrame(delta)          Renderer.cs:1006-1040 -> ptb_compute_frame (async)
    WaitForFences()              Renderer.cs:972       -> ptb_wait
    Render(delta)                Renderer.cs:933-1004  -> ResetFences/ComputeFrame/WaitForFences, then the
                                                          headless present-equivalent instead of DrawFrame
    Dispose()                    Renderer.cs:1192-1216 -> ptb_destroy

Window, surface, swapchain, render pass and graphics pipeline do not exist here (no graphics
stack on the box): `Window` only carries FramebufferSize, and Render() ends in a device buffer.
Failures raise `PtbError` (an Exception with the native message), like the reference's
`throw new Exception(...)` on non-Success results.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

from .tracer import NativeTracer


@dataclass
class Window:
    """Stand-in for Silk.NET's IWindow: the reference creates 1920x1080 (App.cs:27) and the
    renderer only reads Window.FramebufferSize (Renderer.cs:128-129)."""
    FramebufferSize: Tuple[int, int] = (1920, 1080)


class Renderer:
    def __init__(self, window: Optional[Window] = None, *, device: int = 0, keep_f32: bool = False,
                 keep_hit_ids: bool = False, uv_div=None):
        self.Window = window or Window()
        self._device = device
        self._keep_f32 = keep_f32
        self._keep_hit = keep_hit_ids
        self._uv_div = uv_div
        self._tracer: Optional[NativeTracer] = None
        self._pipeline_ready = False
        self._disposed = False
        self._fence_signalled = True       # ComputeFence is created signalled (Renderer.cs:396-402)

    # Renderer.cs:66-84 (compute-relevant subset, same order)
    def Init(self) -> None:
        self.CreateResources()
        self.CreateComputePipeline()

    def CreateResources(self) -> None:
        w, h = self.Window.FramebufferSize
        self._tracer = NativeTracer(w, h, device=self._device, keep_f32=self._keep_f32,
                                    keep_hit_ids=self._keep_hit, uv_div=self._uv_div)

    def CreateComputePipeline(self) -> None:
        if self._tracer is None:
            raise Exception("CreateComputePipeline before CreateResources")
        # The kernel is linked into libptb200.so; nothing to load at run time (the reference
        # reads Assets/Shaders/Compiled/Ray/Test.spirv here, Renderer.cs:356-357).
        self._pipeline_ready = True

    def Update(self, deltaTime: float) -> None:     # Renderer.cs:86-89: empty in the reference
        pass

    def ResetFences(self) -> None:                  # Renderer.cs:970
        self._fence_signalled = False

    def ComputeFrame(self, delta: float) -> None:   # `delta` is unused in the reference too
        if not self._pipeline_ready:
            raise Exception("ComputeFrame before CreateComputePipeline")
        self._tracer.compute_frame()

    def WaitForFences(self) -> None:                # Renderer.cs:972
        self._tracer.wait()
        self._fence_signalled = True

    def Render(self, delta: float) -> None:
        """One frame: compute (synchronous, like the reference) then the present-equivalent."""
        self.ResetFences()
        self.ComputeFrame(delta)
        self.WaitForFences()
        self._tracer.present(srgb=True)

    # Headless consumers (replace the swapchain)
    def ReadImage(self) -> np.ndarray:
        return self._tracer.read_rgba8()

    def ReadPresented(self) -> np.ndarray:
        return self._tracer.read_present()

    @property
    def Tracer(self) -> NativeTracer:
        return self._tracer

    def Dispose(self) -> None:
        if self._disposed:
            return
        if self._tracer is not None:
            self._tracer.destroy()
        self._disposed = True

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.Dispose()
