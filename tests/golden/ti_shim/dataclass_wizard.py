"""Stub of dataclass_wizard for importing the reference (only the base class name is needed)."""


class YAMLWizard:
    pass
