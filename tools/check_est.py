import sys, os, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sklearn.utils.estimator_checks import check_estimator
from dask_ml_b200.cluster import KMeans
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    res = check_estimator(KMeans(), on_fail=None)
bad = [r for r in res if r["status"] not in ("passed", "skipped", "xfail")]
print("total", len(res), "failed", len(bad))
for r in bad:
    print(r["check_name"], "|", str(r["exception"]).splitlines()[0][:160] if r["exception"] else "")
