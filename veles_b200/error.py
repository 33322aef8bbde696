"""Exception hierarchy (reference: veles/error.py)."""


class VelesException(Exception):
    """Base class of all framework errors."""


class Bug(VelesException):
    """Internal invariant violated."""


class BadFormatError(VelesException):
    """Input data has the wrong format."""


class AlreadyExistsError(VelesException):
    pass


class NotExistsError(VelesException):
    pass


class MasterSlaveCommunicationError(VelesException):
    pass


class DeviceNotFoundError(VelesException):
    pass


class BackendNotSupportedError(VelesException):
    """Raised for backends that have no B200 equivalent (e.g. ``ocl``)."""
