"""Support libraries shared by the harness, CLI and the single-box master (reference L2 layer)."""
