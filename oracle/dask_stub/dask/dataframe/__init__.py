"""``dask.dataframe``: only the names the reference uses in ``isinstance`` checks and type aliases."""
from . import utils  # noqa: F401


class DataFrame:
    pass


class Series:
    pass
