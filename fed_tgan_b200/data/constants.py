"""Column-kind tags (reference: ``Server/dtds/data/constants.py:1-3``; the client copy adds
``BIMODAL``, ``Client/.../constants.py:4``).  The *file-level* meta written to JSON uses the
reference's misspelt ``"continous"`` so that existing meta files stay readable."""

CATEGORICAL = "categorical"
CONTINUOUS = "continuous"
ORDINAL = "ordinal"
BIMODAL = "bimodal"

#: spelling used in the file-level meta JSON (``file_generator.py:214`` in the reference)
FILE_META_CONTINUOUS = "continous"
#: placeholder for a missing cell (``file_generator.py:115-116``)
EMPTY_TOKEN = "empty"
