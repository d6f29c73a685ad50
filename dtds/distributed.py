"""``python -m dtds.distributed -ip -port -rank -world_size -epoch[s] -datapath ...``

Same flags and defaults as the reference (``Server/dtds/distributed.py:894-932``); the
implementation is ``fed_tgan_b200.parallel.launcher``.  The reference's public names are
re-exported so code written against ``dtds.distributed`` keeps importing.
"""

from fed_tgan_b200.parallel.client import MDGANClient  # noqa: F401
from fed_tgan_b200.parallel.federator import MDGANServer  # noqa: F401
from fed_tgan_b200.parallel.groups import _average_state_dicts as average_model  # noqa: F401
from fed_tgan_b200.parallel.launcher import main, run  # noqa: F401
from fed_tgan_b200.parallel.weights import softmax  # noqa: F401

if __name__ == "__main__":
    main()
