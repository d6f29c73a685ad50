"""``python similarity_analysis.py -nepoch N`` (reference: ``Server/similarity_analysis.py``)."""
from fed_tgan_b200.eval.similarity import main, stat_sim_normalize  # noqa: F401

if __name__ == "__main__":
    main()
