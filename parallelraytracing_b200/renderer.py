"""Python mirror of the reference's backend interface for the hot path.

  B200Renderer.{Init, ProgressiveRender, SetCamera}   src/core/renderer.h:8-16
  Film.{Clear, AddSampleBufferGPU, UpdateDisplay}      src/core/film.h:22-34, film.cu:25-35, 79-132
  call pattern                                         main.cpp:138-170, 506-532

Same names, argument meaning and error behaviour as the C++ shim in host/b200_renderer.{h,cpp}:
failures in Init are logged and the renderer becomes a no-op (optix/renderer.cpp:1006-1014).
Everything computes on the GPU through the C ABI (capi.Context); there is no CPU path here.
"""
import sys

import numpy as np

from .capi import Context, PrtError


class Film:
    """Sum-of-samples accumulator + per-pixel weight, device resident (Film::m_AccumGPU /
    m_WeightsGPU, film.h:44-61). The weight is uniform over the frame because every
    ProgressiveRender adds whole-frame samples, so it is kept as one scalar."""

    def __init__(self, width, height):
        self.width, self.height = int(width), int(height)
        self._ctx = None

    def GetWidth(self): return self.width
    def GetHeight(self): return self.height

    def _bind(self, ctx):
        self._ctx = ctx

    def Clear(self):                               # film.cu:25-35
        if self._ctx is not None:
            self._ctx.clear_accum()

    def Weight(self):
        return 0 if self._ctx is None else self._ctx.accum_spp()

    def ReadAccum(self, out=None):
        """Raw float sum (H, W, 3): the reference keeps this private (film.h:50); tests need it. `out`: reuse this array
        (e.g. one in page-locked memory, capi.pinned_copy) instead of allocating."""
        return self._ctx.read_accum(out)

    def UpdateDisplay(self, exposure=1.0, gamma=2.2):   # film.cu:101-132 / 134-194
        return self._ctx.resolve_display(exposure, gamma)


class B200Renderer:
    def __init__(self, device=0, max_depth=20, samples_per_call=1, seed=1):
        # device: a CUDA ordinal, or a list of ordinals = ONE renderer over several GPUs of this process (prt_create_multi)
        self.m_Device, self.m_Depth = device, max_depth       # depth 20 = cpu/renderer.h:34
        self.m_SamplesPerCall, self.m_Seed = max(1, samples_per_call), seed
        self.m_NextSample = 0
        self.m_Film = self.m_Scene = self.m_Camera = None
        self.ctx = None
        self.ok = False

    # -- the three methods of the reference interface ------------------------------------
    def Init(self, film, scene, camera):
        self.m_Film, self.m_Scene, self.ok = film, scene, False
        try:
            if self.ctx is None:
                self.ctx = Context(self.m_Device)
            self.ctx.set_scene_build(scene)            # scene upload + BVH build in one call (b200_renderer.cpp does the same)
            self.ok = True
        except PrtError as e:
            print("[B200Renderer]", e, file=sys.stderr)
            return
        self.SetCamera(camera)
        film._bind(self.ctx)

    def SetCamera(self, camera):
        self.m_Camera = camera
        if self.ctx is None or not self.ok:
            return
        try:
            self.ctx.set_camera(camera)
        except PrtError as e:
            print("[B200Renderer]", e, file=sys.stderr)
            self.ok = False

    def ProgressiveRender(self):
        """Adds m_SamplesPerCall (default exactly 1, like every reference backend) samples per
        pixel to the film. Synchronous."""
        if self.ctx is None or not self.ok or self.m_Film is None or self.m_Camera is None:
            return
        if self.m_Film.width != self.m_Camera.width or self.m_Film.height != self.m_Camera.height:
            return                                   # size guard, cuda_wavefront/renderer.cu:374-391
        try:
            self.ctx.render(self.m_SamplesPerCall, self.m_Depth, self.m_Seed, self.m_NextSample)
        except PrtError as e:
            print("[B200Renderer]", e, file=sys.stderr)
            return
        self.m_NextSample += self.m_SamplesPerCall

    # -- extras ---------------------------------------------------------------------------
    def SetMaxDepth(self, d): self.m_Depth = int(d)
    def SetSamplesPerCall(self, k): self.m_SamplesPerCall = max(1, int(k))
    def SetSeed(self, s): self.m_Seed = int(s)
    def SetNextSample(self, s): self.m_NextSample = int(s)

    def Stats(self):
        return self.ctx.stats()

    def close(self):
        if self.ctx is not None:
            self.ctx.close()
            self.ctx = None
