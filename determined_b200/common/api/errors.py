"""REST client errors (reference ``harness/determined/common/api/errors.py``)."""
from typing import Any, Optional


class BadRequestException(Exception):
    def __init__(self, message: str) -> None:
        super().__init__(message)
        self.message = message


class BadResponseException(Exception):
    pass


class MasterNotFoundException(BadRequestException):
    pass


class APIException(BadRequestException):
    """An HTTP error status from the master."""

    def __init__(self, status_code: int, message: str, body: Any = None) -> None:
        super().__init__(message)
        self.status_code = status_code
        self.body = body


class NotFoundException(APIException):
    def __init__(self, message: str, body: Any = None) -> None:
        super().__init__(404, message, body)


class UnauthenticatedException(APIException):
    def __init__(self, message: str = "Unauthenticated", body: Any = None) -> None:
        super().__init__(401, message, body)


class ForbiddenException(APIException):
    def __init__(self, message: str = "Forbidden", body: Any = None) -> None:
        super().__init__(403, message, body)


class CorruptTokenCacheException(Exception):
    pass


class InvalidCredentialsException(Exception):
    pass


class CorruptCertificateCacheException(Exception):
    """The store of accepted master certificates cannot be read."""


class DeleteFailedException(Exception):
    """The master reported that deleting an object failed."""


class EmptyResultException(Exception):
    """A lookup that must find something found nothing."""
