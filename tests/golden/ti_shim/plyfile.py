"""Import stub of `plyfile` -- TEST INFRASTRUCTURE ONLY (the reference's GaussianPointCloudScene
imports PlyData / PlyElement at module level; only its .ply export / import, which the caller tests
do not use, would touch them)."""


class _Unavailable:
    def __init__(self, *a, **k):
        raise RuntimeError("plyfile is not installed here; this is an import stub")

    @classmethod
    def read(cls, *a, **k):
        raise RuntimeError("plyfile is not installed here; this is an import stub")

    @classmethod
    def describe(cls, *a, **k):
        raise RuntimeError("plyfile is not installed here; this is an import stub")


class PlyData(_Unavailable):
    pass


class PlyElement(_Unavailable):
    pass
