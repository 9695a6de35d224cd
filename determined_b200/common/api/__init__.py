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
from determined_b200.common.api._helpers import (  # noqa: E402,F401
    AnyNTSC, NTSC_Kind, PageOpts, WARNING_MESSAGE_MAP, create_group_assignment_request, create_user_assignment_request, get_default_master_url,
    get_ntsc_details, group_name_to_group_id, not_found_errs, read_paginated, role_name_to_role_id, usernames_to_user_ids, wait_for_ntsc_state,
    wait_for_task_ready, workspace_by_name)
from determined_b200.common.api.authentication import salt_and_hash  # noqa: E402,F401
