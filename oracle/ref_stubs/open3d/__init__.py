class _IO:
    @staticmethod
    def read_triangle_mesh(*a, **k):
        return None
io = _IO()
class geometry:  # noqa: N801
    class TriangleMesh:
        pass
