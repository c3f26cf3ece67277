"""Stand-in so that ``dask_ml/_compat.py`` can read ``distributed.__version__``.  TEST INFRASTRUCTURE ONLY."""
__version__ = "2025.5.0"
