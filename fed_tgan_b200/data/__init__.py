from .constants import CATEGORICAL, CONTINUOUS, ORDINAL  # noqa: F401
from .load import prepare_data, encode_data_with_meta_labelencoder, load_datapath  # noqa: F401
from .file_generator import FileGenerator  # noqa: F401
from .transform import Transform  # noqa: F401
from .date import Date  # noqa: F401
