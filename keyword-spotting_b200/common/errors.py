"""Exception hierarchy of the reference, same names and message prefixes
(``/root/reference/kws/common/errors.py:6-32``), so callers that catch the
reference's exceptions keep working when they switch to this package."""

from typing import Any, Optional, Type


class KWSError(Exception):
    def __init__(self, message: str = "An error occurred in KWS"):
        self.message = message
        super().__init__(self.message)


class DatasetError(KWSError):
    def __init__(self, message: str = "Dataset error"):
        super().__init__(f"Dataset error: {message}")


class ModelError(KWSError):
    def __init__(self, message: str = "Model error"):
        super().__init__(f"Model error: {message}")


class AudioProcessingError(KWSError):
    def __init__(self, message: str = "Audio processing error"):
        super().__init__(f"Audio processing error: {message}")


class NativeLibraryError(KWSError):
    """The CUDA library is missing or a kernel call failed.  There is no CPU fallback."""

    def __init__(self, message: str = "native library error"):
        super().__init__(f"kws_b200 native library: {message}")


def handle_error(error: Exception, custom_error: Optional[Type[KWSError]] = None,
                 msg: Optional[str] = None, re_raise: bool = True) -> Any:
    """Same contract as the reference's ``handle_error`` (errors.py:35-63) minus loguru."""
    import logging

    error_msg = msg if msg else str(error)
    logging.getLogger("kws_b200").error("%s: %s", error_msg, error)
    if re_raise:
        if custom_error:
            raise custom_error(error_msg) from error
        raise error
    return None
