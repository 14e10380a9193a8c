"""Stand-in for dataclass_wizard -- TEST INFRASTRUCTURE ONLY.  The reference's config dataclasses
derive from YAMLWizard (GaussianPointCloudRasterisation.py:777, GaussianPointTrainer.py:34); this
class gives them the two calls the reference's entry scripts make, `from_yaml_file` /
`to_yaml_file` (gaussian_point_train.py:12-19), with dataclass_wizard's default key style
(`near-plane` <-> `near_plane`) and nested dataclasses."""
import dataclasses


def _build(cls, data):
    if not dataclasses.is_dataclass(cls) or not isinstance(data, dict):
        return data
    kwargs = {}
    fields = {f.name: f for f in dataclasses.fields(cls)}
    for key, value in data.items():
        name = str(key).replace("-", "_")
        if name not in fields:
            continue
        ftype = fields[name].type
        kwargs[name] = _build(ftype, value) if isinstance(ftype, type) else value
    return cls(**kwargs)


def _dump(obj):
    if dataclasses.is_dataclass(obj) and not isinstance(obj, type):
        return {f.name.replace("_", "-"): _dump(getattr(obj, f.name)) for f in dataclasses.fields(obj)}
    return obj


class YAMLWizard:
    @classmethod
    def from_yaml_file(cls, path):
        import yaml
        with open(path) as f:
            return _build(cls, yaml.safe_load(f) or {})

    def to_yaml_file(self, path):
        import yaml
        with open(path, "w") as f:
            yaml.safe_dump(_dump(self), f)
