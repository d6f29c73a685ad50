"""Stand-alone trainer entry (reference: ``Server/dtds/local.py``)."""
from .models.synthesizer import local_main as main  # noqa: F401

if __name__ == "__main__":
    main()
