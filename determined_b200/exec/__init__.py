"""Process entrypoints run inside an allocation (reference ``harness/determined/exec``)."""
