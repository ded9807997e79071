"""String tables and `create_pipeline` for the command-line callers of the path
(interface of reference stereomatch/cli_common.py:9-52: same keys, same argument order)."""
from .aggregation import Semiglobal
from .cost import SSD, SSDTexture, Birchfield
from .disparity_reduce import DynamicProgramming, WinnerTakesAll
from .pipeline import Pipeline


COST_METHODS = {"ssd": SSD, "ssd-texture": SSDTexture, "birchfield": Birchfield}
AGGREGATION_METHODS = {"sgm": Semiglobal}
DISPARITY_METHODS = {"wta": WinnerTakesAll, "dyn": DynamicProgramming}


def create_pipeline(cost_method: str, disp_method: str, aggr_method: str = None,
                    max_disparity: int = 32) -> Pipeline:
    """cost_method in COST_METHODS, disp_method in DISPARITY_METHODS, aggr_method in
    AGGREGATION_METHODS or anything else (None, "") for no aggregation -- as the reference,
    an unknown aggregation name silently means none; unknown cost/disparity names raise
    KeyError."""
    aggregation = AGGREGATION_METHODS.get(aggr_method, None)
    if aggregation is not None:
        aggregation = aggregation()
    return Pipeline(COST_METHODS[cost_method](max_disparity), DISPARITY_METHODS[disp_method](),
                    aggregation=aggregation)
