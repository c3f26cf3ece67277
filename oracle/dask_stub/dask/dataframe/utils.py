def assert_eq(a, b, **kwargs):  # pragma: no cover - not on the k-means path
    raise NotImplementedError("dask.dataframe is not part of the stub")
