def __getattr__(name):
    def unavailable(*a, **k):
        raise RuntimeError(f"pytorch3d.transforms.{name}: pytorch3d is not installed here (import stub)")
    return unavailable
