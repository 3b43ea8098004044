"""Headless replacement of the progressive viewer (visualization/inspector.h:6-8, inspector.cpp:71-282) and of the
`render` callback of samples/sample1.cpp:447-548.

The reference opens a GLFW/OpenGL window, calls the user's callback once per displayed frame and shows the RGBA8
image it returns; the callback traces one batch, reads the image back and advances `RTProp.totalSamples` (or resets
it when the ImGui panel changed the camera / light / debug mode). There is no graphics stack on a compute box
(window, GL upload: unavailable on this box), so `renderLoop` here runs the same callback protocol for a fixed number
of frames and writes the returned images to PNG files; everything else -- the accumulation protocol, the edits the
panel could make -- is kept call for call.
"""
import numpy as np

from . import rd
from .scenes import CAMERA_DTYPE, RTPROP_DTYPE, SCENEPROP_DTYPE


def write_png(path, rgba):
    from PIL import Image
    Image.fromarray(np.ascontiguousarray(rgba, np.uint8), "RGBA").save(path, format="PNG")


def renderLoop(callback, callbackData, frames=1, out_pattern=None, every=1):
    """`int renderLoop(Callback, void*)` (inspector.h:8) without the window: `callback(callbackData)` must return
    `(image, width, height)` with image an (H, W, 4) uint8 array -- the three out-parameters of the C callback.
    Frame k is written to `out_pattern % k` when `out_pattern` is given and k % every == 0 (always the last one).
    Returns 0 like the reference's loop on a clean exit."""
    for k in range(frames):
        image, width, height = callback(callbackData)
        if image is None or image.shape[0] != height or image.shape[1] != width:
            raise rd.RadianceError("renderLoop: the callback returned no image or a wrong extent")
        if out_pattern is not None and (k % every == 0 or k == frames - 1):
            write_png(out_pattern % k, image)
    return 0


class CbData:
    """The callback data of samples/sample1.cpp:344-361: platform, extent, the handles the panel edits."""

    def __init__(self, plt, extent, rdImage, rdRTProp, rdCamData, rdSceneData):
        self.plt, self.extent = plt, (int(extent[0]), int(extent[1]))
        self.rdImage, self.rdRTProp, self.rdCamData, self.rdSceneData = rdImage, rdRTProp, rdCamData, rdSceneData
        self.imageSize = self.extent[0] * self.extent[1] * 4
        self.image = np.zeros((self.extent[1], self.extent[0], 4), np.uint8)
        self.updated = False          # set by the edit functions below, consumed by render()
        self.frames = 0

    # ---- what RenderSceneConfigUI (sample1.cpp:501-548) can change; each edit restarts the accumulation ----
    def _edit(self, handle, dtype, fn):
        v = np.zeros(1, dtype)
        rd.ReadBuffer(self.plt, handle, v.nbytes, v)
        fn(v[0])
        rd.WriteBuffer(self.plt, handle, v.nbytes, v)
        self.updated = True

    def set_camera(self, position=None, rotation=None, focal=None):
        def fn(c):
            if position is not None:
                c["x"], c["y"], c["z"] = position
            if rotation is not None:
                c["wx"], c["wy"], c["wz"] = rotation
            if focal is not None:
                c["focalLength"] = focal
        self._edit(self.rdCamData, CAMERA_DTYPE, fn)

    def set_light(self, direction=None, color=None, index=0):
        def fn(s):
            if direction is not None:
                s["lights"][index]["direction"][:len(direction)] = direction
            if color is not None:
                s["lights"][index]["color"][:len(color)] = color
        self._edit(self.rdSceneData, SCENEPROP_DTYPE, fn)

    def set_debug(self, mode):
        self._edit(self.rdRTProp, RTPROP_DTYPE, lambda r: r.__setitem__("debug", mode))

    def total_samples(self):
        v = np.zeros(1, RTPROP_DTYPE)
        rd.ReadBuffer(self.plt, self.rdRTProp, v.nbytes, v)
        return int(v[0]["totalSamples"])


def render(d):
    """The `render` callback of samples/sample1.cpp:447-499: trace one batch, fetch the image, then
    `totalSamples = updated ? 0 : totalSamples + batchSize`. As in the reference, the batch traced in the frame in
    which an edit arrived still accumulates onto the old samples; the restart takes effect from the next frame."""
    updated, d.updated = d.updated, False
    rd.TraceRays(d.plt, 0, 0, 0, d.extent[0], d.extent[1])
    rd.ReadBuffer(d.plt, d.rdImage, d.imageSize, d.image)
    rt = np.zeros(1, RTPROP_DTYPE)
    rd.ReadBuffer(d.plt, d.rdRTProp, rt.nbytes, rt)
    rt[0]["totalSamples"] = 0 if updated else rt[0]["totalSamples"] + rt[0]["batchSize"]
    rd.WriteBuffer(d.plt, d.rdRTProp, rt.nbytes, rt)
    d.frames += 1
    return d.image, d.extent[0], d.extent[1]
