"""REST client (reference ``harness/determined/common/api``)."""
from determined_b200.common.api import authentication, errors  # noqa: F401
from determined_b200.common.api._session import (  # noqa: F401
    BaseSession, Session, UnauthSession, canonicalize_master_url,
)
from determined_b200.common.api.errors import (  # noqa: F401
    APIException, BadRequestException, BadResponseException, ForbiddenException, MasterNotFoundException,
    NotFoundException, UnauthenticatedException,
)
