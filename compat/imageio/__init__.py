"""Minimal OpenCV-backed stand-in for the three imageio calls the reference's demos make (see compat/README.md)."""
import cv2
import numpy as np


def imread(path):
    img = cv2.imread(str(path), cv2.IMREAD_UNCHANGED)
    if img is None:
        raise FileNotFoundError(path)
    return img[..., ::-1] if img.ndim == 3 and img.shape[2] >= 3 else img


def imwrite(path, img):
    a = np.asarray(img)
    cv2.imwrite(str(path), a[..., ::-1] if a.ndim == 3 and a.shape[2] >= 3 else a)


class _Writer:
    def __init__(self, path, fps=25, **_):
        self.path, self.fps, self.vw = str(path), float(fps), None

    def append_data(self, frame):
        a = np.asarray(frame)
        if a.ndim == 2:
            a = np.stack([a] * 3, axis=2)
        a = np.ascontiguousarray(a[..., :3][..., ::-1]).astype(np.uint8)
        if self.vw is None:
            h, w = a.shape[:2]
            self.vw = cv2.VideoWriter(self.path, cv2.VideoWriter_fourcc(*"mp4v"), self.fps, (w, h))
        self.vw.write(a)

    def close(self):
        if self.vw is not None:
            self.vw.release()
            self.vw = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def get_writer(path, fps=25, **kw):
    return _Writer(path, fps=fps, **kw)
