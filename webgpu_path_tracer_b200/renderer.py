"""Host-side mirror of the reference's `Renderer` (/root/reference/renderer.js) over libwgpt.so.

Same method names, argument meaning and per-frame sequence: `initBuffers` uploads the eight
scene arrays + a zeroed framebuffer (renderer.js:68-103), `renderAnimation` runs ONE iteration
of the reference's loop -- frameNum += 1, uniforms {frameNum, resetBuffer, viewMatrix} written,
compute pass of W*H/64 + 1 workgroups, then the present pass (renderer.js:163-191) -- and returns
instead of rescheduling itself (no requestAnimationFrame headless).  The swap chain is a host
array: `last_image` holds what `fs` would have put on the canvas.
"""
import numpy as np

from .api import PathTracer, MODE_EXACT, MODE_FAST
from . import imageio


class Renderer:
    def __init__(self, mode=MODE_FAST, device=0, num_samples=1, max_bounces=100, stratify=False,
                 importance_sampling=False, stack_size=20, **shard):
        self._ctx_args = dict(mode=mode, device=device, num_samples=num_samples, max_bounces=max_bounces,
                              stratify=stratify, importance_sampling=importance_sampling, stack_size=stack_size)
        self._ctx_args.update(shard)
        self.webGPU = None          # the wgpt context takes the place of the WebGPU device wrapper
        self.frameNum = 0.0
        self.last_image = None

    @staticmethod
    def create(**kw):
        """renderer.js:12-16 (device acquisition happens in initBuffers, when the size is known)."""
        return Renderer(**kw)

    def initBuffers(self, scene, camera, WIDTH, HEIGHT):
        """renderer.js:68-103."""
        self.WIDTH, self.HEIGHT = int(WIDTH), int(HEIGHT)
        self.uniforms = {"screenDims": [WIDTH, HEIGHT], "frameNum": 0, "resetBuffer": 0, "viewMatrix": camera.viewMatrix}
        scene.init_mesh_data()
        scene.create_meshes()
        scene.create_bvh()
        self.webGPU = PathTracer(self.WIDTH, self.HEIGHT, **self._ctx_args)
        self.webGPU.write_scene(scene.buffers())
        self.webGPU.write_buffer(3, np.zeros(self.WIDTH * self.HEIGHT * 4, np.float32))
        self.camera = camera

    def setRenderParameters(self, renderParams, camera, WIDTH, HEIGHT):
        """renderer.js:141-160."""
        self.renderParams = renderParams
        self.frameNum = 0.0
        self.camera = camera
        self.WIDTH, self.HEIGHT = int(WIDTH), int(HEIGHT)

    def renderAnimation(self, present=True):
        """One pass of renderer.js:163-191."""
        self.frameNum += 1.0
        moving = bool(self.camera.MOVING or self.camera.keyPress)
        self.uniforms["frameNum"] = self.frameNum
        self.uniforms["resetBuffer"] = 1 if moving else 0
        if moving:
            self.frameNum = 1
            self.camera.keyPress = 0
        self.uniforms["viewMatrix"] = self.camera.viewMatrix
        self.webGPU.write_uniforms(self.uniforms["frameNum"], self.uniforms["resetBuffer"], self.uniforms["viewMatrix"])
        workGroupsNeeded = (self.WIDTH * self.HEIGHT) / 64
        self.webGPU.dispatch(int(workGroupsNeeded + 1))
        if present:
            self.last_image = self.webGPU.present().reshape(self.HEIGHT, self.WIDTH, 4)
        return self.last_image

    def renderFrames(self, n_frames):
        """Batched renderAnimation without camera motion and without per-frame present
        (wgpt_render): frames frameNum+1 .. frameNum+n."""
        first = int(self.frameNum) + 1
        self.webGPU.write_uniforms(first, 0, self.camera.viewMatrix)
        self.webGPU.render(first, int(n_frames))
        self.frameNum += n_frames

    def renderProgressive(self, n_frames, refresh_every, on_refresh):
        """The progressive viewer loop of renderAnimation (renderer.js:163-214: the canvas refreshes while samples
        accumulate) without a canvas: frames are dispatched one by one -- wgpt_dispatch enqueues and returns, the frames
        join one living wavefront -- and every `refresh_every` frames `on_refresh(frameNum, image)` receives what `fs`
        would show (float RGBA, H x W x 4).  Returns the last image."""
        img = None
        for k in range(int(n_frames)):
            self.renderAnimation(present=False)
            if (k + 1) % int(refresh_every) == 0 or k + 1 == int(n_frames):
                img = self.webGPU.present().reshape(self.HEIGHT, self.WIDTH, 4)
                self.last_image = img
                if on_refresh is not None:
                    on_refresh(self.frameNum, img)
        return img

    def present(self, rgba8=False):
        self.webGPU.write_uniforms(self.frameNum, 0, self.camera.viewMatrix)
        return self.webGPU.present(rgba8=rgba8)

    def save_image(self, path):
        """Write what the canvas would show (fs output, RGBA8) to a .png or .ppm file; .pfm gets the float image."""
        if str(path).lower().endswith(".pfm"):
            img = np.asarray(self.present()).reshape(self.HEIGHT, self.WIDTH, 4)
            imageio.write_pfm(path, img)
            return img
        rgba8 = np.asarray(self.present(rgba8="only")).reshape(self.HEIGHT, self.WIDTH, 4)
        (imageio.write_ppm if str(path).lower().endswith(".ppm") else imageio.write_png)(path, rgba8)
        return rgba8
