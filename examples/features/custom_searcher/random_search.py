"""A client-side custom SearchMethod driven by LocalSearchRunner (reference ``examples/custom_search_method``)."""
import pathlib
import random
import sys
import uuid
from typing import List

from determined_b200 import searcher


class RandomSearch(searcher.SearchMethod):
    def __init__(self, max_trials: int, max_length: int, seed: int = 0) -> None:
        self.max_trials, self.max_length = max_trials, max_length
        self.rng = random.Random(seed)
        self.created = self.closed = 0

    def _new(self) -> List[searcher.Operation]:
        self.created += 1
        rid = uuid.uuid4()
        hp = {"learning_rate": 10 ** self.rng.uniform(-3, 0), "global_batch_size": 64, "n_filters1": 32, "n_filters2": 64, "dropout1": 0.25, "dropout2": 0.5}
        return [searcher.Create(rid, hp, None), searcher.ValidateAfter(rid, self.max_length), searcher.Close(rid)]

    def initial_operations(self, st):
        return self._new() + self._new()

    def on_trial_created(self, st, rid):
        return []

    def on_validation_completed(self, st, rid, metric, train_length):
        return []

    def on_trial_closed(self, st, rid):
        self.closed += 1
        if self.created < self.max_trials:
            return self._new()
        return [searcher.Shutdown()] if self.closed == self.max_trials else []

    def on_trial_exited_early(self, st, rid, reason):
        return self.on_trial_closed(st, rid)

    def progress(self, st):
        return self.closed / self.max_trials


if __name__ == "__main__":
    model_dir = pathlib.Path(__file__).resolve().parents[2] / "tutorials" / "mnist_pytorch"
    cfg = {"name": "custom-random", "entrypoint": "model_def:MNistTrial", "searcher": {"name": "custom", "metric": "validation_loss", "unit": "batches"},
           "hyperparameters": {"global_batch_size": 64}}
    runner = searcher.LocalSearchRunner(RandomSearch(max_trials=4, max_length=100), searcher_dir=pathlib.Path("search_state"))
    sys.exit(0 if runner.run(cfg, model_dir=str(model_dir)) else 1)
