"""veles_b200 — a Blackwell (B200, sm_100a) native deep-learning platform with the
capabilities, public API and checkpoint format of Veles (reference: devbib/veles)."""
__version__ = "0.1.0"


def __getattr__(name):
    if name == "Trainer":
        from .trainer import Trainer
        return Trainer
    raise AttributeError(name)
