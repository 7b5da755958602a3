"""Stand-in for the reference's `renderutils.SoftRenderer` (CUDA soft rasteriser, out of the solver hot path).
The constructor takes the reference's keyword arguments (renderutils/softrenderer.py:24-120); frames are produced
headlessly by `src.LBM.utils.save_rendered_image`, which does not call the renderer."""


class SoftRenderer:
    def __init__(self, *args, **kwargs):
        self.args, self.kwargs = args, kwargs
        for k, v in kwargs.items():
            setattr(self, k, v)

    def forward(self, *args, **kwargs):
        raise NotImplementedError("compat SoftRenderer: rendering is done by save_rendered_image (headless projection)")

    __call__ = forward
