"""Launch layers (reference ``harness/determined/launch``).  On one box all three reference
launchers (torch_distributed / deepspeed / horovod) reduce to the same thing — start
``slots_per_trial`` local ranks with the torchrun environment contract — so they share
:func:`determined_b200.launch._common.launch`, which prefers the native ``detb200-launch`` binary
(spawner + pidfd watchdog + ``[rank=N]`` prefixer in one process) and falls back to
``torch.distributed.run`` wrapped by the Python PID server."""
