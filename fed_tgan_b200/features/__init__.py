from .gmm import GMMParams, fit_vbgmm_1d, fit_sklearn_1d  # noqa: F401
from .layout import TableLayout  # noqa: F401
from .transformers import (  # noqa: F401
    Transformer, BGM_CTGAN_Transformer, BGMTransformer, DiscretizeTransformer, GeneralTransformer,
    GMMTransformer, TableganTransformer, decode_train_data,
)
