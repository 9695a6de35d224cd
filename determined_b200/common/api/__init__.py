"""REST client (reference ``harness/determined/common/api``)."""
from typing import Any
from determined_b200.common.api import authentication, errors  # noqa: F401
from determined_b200.common.api._session import (  # noqa: F401
    BaseSession, Session, UnauthSession, canonicalize_master_url,
)
from determined_b200.common.api.errors import (  # noqa: F401
    APIException, BadRequestException, BadResponseException, ForbiddenException, MasterNotFoundException,
    NotFoundException, UnauthenticatedException,
)
from determined_b200.common.api import logs  # noqa: E402,F401
from determined_b200.common.api.logs import pprint_logs, task_logs, trial_logs  # noqa: E402,F401


def get_validation_metric(metric_name: str, validation: Any) -> Any:
    """The named metric out of a validation report in any of the shapes the API has used (reference ``api/metric.py``)."""
    if not validation:
        return None
    m = validation.get("metrics", validation) if isinstance(validation, dict) else getattr(validation, "metrics", None)
    if isinstance(m, dict):
        m = m.get("avgMetrics", m.get("validation_metrics", m))
    return (m or {}).get(metric_name) if isinstance(m, dict) else None
