"""Image preprocessing pipeline shared by the image loaders.

Parity target: ``veles/loader/image.py:106-806``: colour space conversion, optional Sobel channel, scaling
(optionally keeping the aspect ratio onto a background), cropping (centre / random, optionally around a
bbox), mirroring (fixed / random), rotations, ``samples_inflation`` = #rotations x #mirror states x
crop_number (``:406-558``).  OpenCV does the pixel work; output is HWC float32/uint8 (NHWC batches)."""
from __future__ import annotations

import numpy

try:
    import cv2
except ImportError:  # pragma: no cover
    cv2 = None

COLOR_CODES = {}
if cv2 is not None:
    COLOR_CODES = {("BGR", "RGB"): cv2.COLOR_BGR2RGB, ("BGR", "GRAY"): cv2.COLOR_BGR2GRAY,
                   ("BGR", "HSV"): cv2.COLOR_BGR2HSV, ("BGR", "YCR_CB"): cv2.COLOR_BGR2YCrCb,
                   ("RGB", "BGR"): cv2.COLOR_RGB2BGR, ("RGB", "GRAY"): cv2.COLOR_RGB2GRAY,
                   ("RGB", "HSV"): cv2.COLOR_RGB2HSV, ("RGB", "YCR_CB"): cv2.COLOR_RGB2YCrCb,
                   ("GRAY", "BGR"): cv2.COLOR_GRAY2BGR, ("GRAY", "RGB"): cv2.COLOR_GRAY2RGB}


class ImagePipeline(object):
    """kwargs: color_space, source color_space ("BGR" for cv2.imread), size/scale, scale_maintain_aspect_ratio,
    background_color, crop (h, w) absolute or fractional, crop_number, mirror (False|True|"random"),
    rotations (radians), add_sobel, smart_crop."""

    def __init__(self, **kwargs):
        self.color_space = kwargs.get("color_space", "RGB")
        self.source_color_space = kwargs.get("source_color_space", "BGR")
        self.scale = kwargs.get("scale", kwargs.get("size", 1.0))   # float factor or (h, w)
        self.scale_maintain_aspect_ratio = kwargs.get("scale_maintain_aspect_ratio", False)
        self.background_color = kwargs.get("background_color", 0)
        self.crop = kwargs.get("crop", None)
        self.crop_number = int(kwargs.get("crop_number", 1))
        self.mirror = kwargs.get("mirror", False)
        self.rotations = tuple(kwargs.get("rotations", (0.0,)))
        self.add_sobel = kwargs.get("add_sobel", False)
        self.rand = kwargs.get("prng", None)

    @property
    def samples_inflation(self):
        return (2 if self.mirror is True else 1) * len(self.rotations) * self.crop_number

    def decode(self, path):
        if cv2 is None:
            raise RuntimeError("OpenCV (cv2) is required by the image loaders")
        img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        if img is None:
            raise ValueError("Cannot decode the image %s" % path)
        return img

    def convert_color(self, img):
        src = self.source_color_space if img.ndim == 3 else "GRAY"
        if src == self.color_space:
            return img
        code = COLOR_CODES.get((src, self.color_space))
        if code is None:
            raise ValueError("Unsupported colour conversion %s -> %s" % (src, self.color_space))
        return cv2.cvtColor(img, code)

    def scale_image(self, img):
        if isinstance(self.scale, (int, float)):
            if self.scale == 1.0:
                return img
            h, w = int(round(img.shape[0] * self.scale)), int(round(img.shape[1] * self.scale))
        else:
            h, w = self.scale
        if not self.scale_maintain_aspect_ratio:
            return cv2.resize(img, (w, h), interpolation=cv2.INTER_AREA if h < img.shape[0] else cv2.INTER_CUBIC)
        ratio = min(h / img.shape[0], w / img.shape[1])
        nh, nw = max(1, int(img.shape[0] * ratio)), max(1, int(img.shape[1] * ratio))
        small = cv2.resize(img, (nw, nh), interpolation=cv2.INTER_AREA)
        canvas = numpy.full((h, w) + img.shape[2:], self.background_color, dtype=img.dtype)
        y0, x0 = (h - nh) // 2, (w - nw) // 2
        canvas[y0:y0 + nh, x0:x0 + nw] = small
        return canvas

    def crop_image(self, img, random=False):
        if self.crop is None:
            return img
        ch, cw = self.crop
        if isinstance(ch, float) and ch <= 1.0:
            ch, cw = int(img.shape[0] * ch), int(img.shape[1] * cw)
        if ch > img.shape[0] or cw > img.shape[1]:
            raise ValueError("Crop %s is larger than the image %s" % ((ch, cw), img.shape[:2]))
        if random and self.rand is not None:
            y0 = int(self.rand.randint(0, img.shape[0] - ch + 1))
            x0 = int(self.rand.randint(0, img.shape[1] - cw + 1))
        else:
            y0, x0 = (img.shape[0] - ch) // 2, (img.shape[1] - cw) // 2
        return img[y0:y0 + ch, x0:x0 + cw]

    def rotate(self, img, angle):
        if not angle:
            return img
        h, w = img.shape[:2]
        m = cv2.getRotationMatrix2D((w / 2.0, h / 2.0), numpy.degrees(angle), 1.0)
        return cv2.warpAffine(img, m, (w, h), borderValue=self.background_color)

    def sobel(self, img):
        gray = img if img.ndim == 2 else cv2.cvtColor(img, cv2.COLOR_RGB2GRAY)
        gx = cv2.Sobel(gray, cv2.CV_32F, 1, 0, ksize=3)
        gy = cv2.Sobel(gray, cv2.CV_32F, 0, 1, ksize=3)
        mag = numpy.sqrt(gx * gx + gy * gy)
        mx = mag.max()
        mag = (mag * (255.0 / mx) if mx > 0 else mag).astype(img.dtype)
        if img.ndim == 2:
            img = img[..., None]
        return numpy.concatenate([img, mag[..., None]], axis=2)

    def process(self, img, training=False):
        """Yield every inflated variant of one decoded image as an HWC array."""
        img = self.scale_image(self.convert_color(img))
        mirrors = (False, True) if self.mirror is True else (False,)
        for angle in self.rotations:
            rot = self.rotate(img, angle)
            for m in mirrors:
                mir = rot[:, ::-1] if m else rot
                if self.mirror == "random" and training and self.rand is not None and self.rand.randint(0, 2):
                    mir = mir[:, ::-1]
                for _ in range(self.crop_number):
                    out = self.crop_image(mir, random=training and self.crop_number >= 1 and self.rand is not None)
                    if self.add_sobel:
                        out = self.sobel(out)
                    if out.ndim == 2:
                        out = out[..., None]
                    yield numpy.ascontiguousarray(out)

    def output_shape(self, sample_image):
        return next(self.process(sample_image)).shape
