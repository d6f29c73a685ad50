from .table import DeviceTable, cumulative_tables  # noqa: F401
from .step import TrainEngine  # noqa: F401
