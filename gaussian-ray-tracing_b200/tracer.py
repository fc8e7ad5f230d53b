"""Python mirror of the reference host API for the hot path (same names and call order as
src/GaussianTracer.h:27-111, src/Camera.h:9-64, src/CUDAOutputBuffer.h:12-40), driving libgrt.so.

    tracer = GaussianTracer(ply_path_or_records)   # ctor loads the PLY            (GaussianTracer.cpp:20-52)
    tracer.setSize(w, h)                           # before initializeOptix        (main.cpp:68)
    tracer.initializeOptix()                       # name kept; builds the LBVH    (GaussianTracer.cpp:72-83)
    tracer.updateCamera(camera, True); tracer.render(output_buffer)                (main.cpp:92-100)

The C++ twin with identical semantics is host/GaussianTracer.{h,cpp}.  No arithmetic of the path lives here.
"""
from __future__ import annotations

import numpy as np

from . import scene as _scene


class Camera:
    """src/Camera.h:9-64 (look-at camera; UVWFrame in src/Camera.cpp:3-13)."""

    def __init__(self):
        self.m_eye = np.zeros(3, np.float32)
        self.m_lookat = np.zeros(3, np.float32)
        self.m_up = np.zeros(3, np.float32)
        self.m_fovY = 0.0
        self.m_aspectRatio = 1.0

    def eye(self): return self.m_eye
    def lookat(self): return self.m_lookat
    def setEye(self, v): self.m_eye = np.asarray(v, np.float32)
    def setLookat(self, v): self.m_lookat = np.asarray(v, np.float32)
    def setUp(self, v): self.m_up = np.asarray(v, np.float32)
    def setFovY(self, v): self.m_fovY = float(v)
    def setAspectRatio(self, v): self.m_aspectRatio = float(v)

    def UVWFrame(self):
        return _scene.camera_uvw(self.m_eye, self.m_lookat, self.m_up, self.m_fovY, self.m_aspectRatio)


class CUDAOutputBuffer:
    """Headless stand-in for src/CUDAOutputBuffer.h:12-40: a plain device buffer instead of a GL PBO
    (no graphics stack on the box).  map() returns the device uchar3*; getPBO() is dropped."""

    def __init__(self, ctx, width, height):
        self.ctx = ctx
        self.m_width = self.m_height = 0
        self.d_ptr = 0
        self.stream = None
        self.resize(width, height)

    def resize(self, width, height):
        if self.d_ptr:
            self.ctx.free(self.d_ptr)
        self.m_width, self.m_height = int(width), int(height)
        self.d_ptr = self.ctx.malloc(max(1, self.m_width * self.m_height * 3))

    def setStream(self, stream): self.stream = stream
    def width(self): return self.m_width
    def height(self): return self.m_height
    def map(self): return self.d_ptr
    def unmap(self): pass

    def download(self) -> np.ndarray:
        return self.ctx.to_host(self.d_ptr, (self.m_height, self.m_width, 3), np.uint8)


class GaussianTracer:
    """Same surface as the reference class; `params` is public and re-read on every render()."""

    def __init__(self, source, device: int = 0, ctx=None):
        from . import Context, GrtParams
        if isinstance(source, str):
            names, data = _scene.read_ply_properties(source)
            col = [names.index(n) for n in _scene.PLY_PROPERTIES]  # missing property -> ValueError (happly throws too)
            self.particles = _scene.activate(data[:, col])
        else:
            self.particles = np.ascontiguousarray(source, dtype=np.float32)
        self.ctx = ctx or Context(device)
        self.alpha_min = 0.01                       # GaussianTracer.cpp:45
        self.params = GrtParams()
        self.stream = None
        self.primitives = []                        # list of primitive dicts (Primitives::m_primitives)
        self.current_lookat = np.zeros(3, np.float32)
        self._initialized = False

    # -- reference API
    def setSize(self, width, height):
        self.params.width, self.params.height = int(width), int(height)

    def getGaussianCenter(self):
        from .scene import gaussian_center
        return gaussian_center(self.particles)

    def initializeOptix(self):
        """Keeps the reference name; uploads the particles and builds the LBVH on the device."""
        self.ctx.set_particles(self.particles, self.alpha_min)
        handle, d_particles = self.ctx.build_particle_accel(self.stream)
        p = self.params
        p.output_buffer = None                      # initializeParams, GaussianTracer.cpp:475-486
        p.handle, p.d_particles = handle, d_particles
        p.t_min, p.t_max, p.minTransmittance, p.alpha_min = 1e-3, 1e5, 0.001, self.alpha_min
        p.sh_degree_max = 0
        p.mesh_handle, p.d_meshes = 0, None
        p.mode_fisheye = 0
        p.type = 0
        p.traceState = None
        self._initialized = True

    def updateCamera(self, camera: Camera, camera_changed: bool = True):
        if not camera_changed:
            return False
        camera.setAspectRatio(float(np.float32(self.params.width) / np.float32(self.params.height)))
        U, V, W = camera.UVWFrame()
        for k in range(3):
            self.params.eye[k] = float(camera.eye()[k])
            self.params.U[k], self.params.V[k], self.params.W[k] = float(U[k]), float(V[k]), float(W[k])
        self.current_lookat = np.asarray(camera.lookat(), np.float32)
        return False

    def render(self, output_buffer: CUDAOutputBuffer):
        if not self._initialized:
            raise RuntimeError("initializeOptix() must be called before render()")
        if self.params.mode_fisheye:                # GaussianTracer.cpp:510-513
            self.ctx.memset(output_buffer.map(), 0, output_buffer.width() * output_buffer.height() * 3)
        self.params.output_buffer = output_buffer.map()
        self.ctx.launch(self.params, self.stream)
        output_buffer.unmap()
        self.ctx.sync(self.stream)                  # CUDA_SYNC_CHECK, GaussianTracer.cpp:537

    def setRenderType(self, t): self.params.type = int(t)
    def getPrimitives(self): return self.primitives

    def _primitive_position(self):                  # GaussianTracer.cpp:580-588
        return _scene.primitive_position(np.array(list(self.params.eye), np.float32), self.current_lookat)

    def _rebuild_meshes(self):
        self.ctx.set_meshes(self.primitives)
        self.params.mesh_handle, dm = self.ctx.build_mesh_accel(self.stream)
        self.params.d_meshes = dm

    def createPlane(self):
        p = _scene.create_plane(self._primitive_position())
        p["instanceIndex"] = len(self.primitives)
        self.primitives.append(p)
        self._rebuild_meshes()
        return p

    def createSphere(self):
        p = _scene.create_sphere(self._primitive_position())
        p["instanceIndex"] = len(self.primitives)
        self.primitives.append(p)
        self._rebuild_meshes()
        return p

    def createLoadMesh(self, filename):
        raise NotImplementedError("OBJ loading is out of scope (tinyobjloader submodule absent in the reference, SURVEY 2.1 #21)")

    def updateInstanceTransforms(self, p):          # GaussianTracer.cpp:711-736
        self.primitives[p["instanceIndex"]] = p
        self._rebuild_meshes()

    def removePrimitive(self):                      # GaussianTracer.cpp:553-566
        self.primitives = []
        self._rebuild_meshes()
