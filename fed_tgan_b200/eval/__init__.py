from .similarity import stat_sim_normalize, column_similarity, analyse_run  # noqa: F401
from .utility import real_res, utility_difference  # noqa: F401
