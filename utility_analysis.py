"""``python utility_analysis.py -train_path -test_path -synthetic_path``
(reference: ``Server/utility_analysis.py``)."""
from fed_tgan_b200.eval.utility import main, real_res  # noqa: F401

if __name__ == "__main__":
    main()
