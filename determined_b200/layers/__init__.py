from determined_b200.layers._workload_sequencer import WorkloadSequencer, make_compatibility_workloads  # noqa: F401
