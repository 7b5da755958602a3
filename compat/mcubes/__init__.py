"""Stand-in for PyMCubes (not installed here); see compat/README.md."""


def marching_cubes(*args, **kwargs):
    raise NotImplementedError("PyMCubes is not available in this image (viz only, outside the solver hot path)")
