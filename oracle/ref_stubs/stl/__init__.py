class _M:  # numpy-stl stand-in; meshes are visualisation-only and out of scope
    pass
mesh = _M()
