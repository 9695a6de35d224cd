"""``core_context.experimental`` (reference ``harness/determined/core/_experimental.py:1-77``):
links an inference/fine-tuning trial to the checkpoint or model version it consumed."""
import logging
from typing import Any, Optional

logger = logging.getLogger("determined.core")


class ExperimentalCoreContext:
    def __init__(self, session: Any, trial_id: int) -> None:
        self._session, self._trial_id = session, trial_id

    def report_task_using_checkpoint(self, checkpoint: Any) -> None:
        uuid = getattr(checkpoint, "uuid", checkpoint)
        self._session.post("/api/v1/trial-source-info", json={"trialSourceInfo": {
            "checkpointUuid": uuid, "trialId": self._trial_id,
            "trialSourceInfoType": "TRIAL_SOURCE_INFO_TYPE_INFERENCE"}})

    def report_task_using_model_version(self, model_version: Any) -> None:
        ckpt = getattr(model_version, "checkpoint", None)
        self._session.post("/api/v1/trial-source-info", json={"trialSourceInfo": {
            "checkpointUuid": getattr(ckpt, "uuid", None), "trialId": self._trial_id,
            "trialSourceInfoType": "TRIAL_SOURCE_INFO_TYPE_INFERENCE",
            "modelId": getattr(model_version, "model_id", None),
            "modelVersion": getattr(model_version, "model_version", None)}})


class DummyExperimentalCoreContext(ExperimentalCoreContext):
    def __init__(self) -> None:
        pass

    def report_task_using_checkpoint(self, checkpoint: Any) -> None:
        logger.info("report_task_using_checkpoint(%s)", getattr(checkpoint, "uuid", checkpoint))

    def report_task_using_model_version(self, model_version: Any) -> None:
        logger.info("report_task_using_model_version(%s)", model_version)
