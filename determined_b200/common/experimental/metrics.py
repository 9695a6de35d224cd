from determined_b200.experimental.client import TrainingMetrics, TrialMetrics, ValidationMetrics  # noqa: F401
