"""Metaclass mapping ``MAPPING = "name"`` class attributes to classes
(reference: veles/mapped_object_registry.py:36-54).  Used for loaders, normalizers,
snapshotters, publishing backends: ``Registry.<mapping>[name] -> class``."""


class MappedObjectsRegistry(type):
    mapping = "You must define \"mapping\" static attribute in your metaclass"
    base = object

    def __init__(cls, name, bases, clsdict):
        yours = set(cls.mro())
        mine = set(cls.base.mro()) if isinstance(cls.base, type) else set()
        left = yours - mine
        mapping = clsdict.get("MAPPING", None)
        registry = getattr(type(cls), type(cls).mapping, None)
        if registry is None:
            registry = {}
            setattr(type(cls), type(cls).mapping, registry)
        if len(left) > 1 and mapping is not None:
            registry[mapping] = cls
        super(MappedObjectsRegistry, cls).__init__(name, bases, clsdict)
