from determined_b200.master.searcher.methods import (  # noqa: F401
    ASHASearch, Context, CustomSearch, GridSearch, RandomSearch, SearchMethod, TournamentSearch,
    adaptive_brackets, bracket_max_concurrent, bracket_max_trials, length_units, make_adaptive_asha,
    new_search_method,
)
from determined_b200.master.searcher.operations import (  # noqa: F401
    Close, Create, ExitedReason, Shutdown, ValidateAfter, new_create, op_from_json,
)
from determined_b200.master.searcher.searcher import Searcher, simulate  # noqa: F401
