"""Agent daemon (reference ``agent/``): ``python -m determined_b200.agent --master-url ...``."""
from determined_b200.agent.agent import Agent, ProcessRuntime, detect_slots  # noqa: F401
