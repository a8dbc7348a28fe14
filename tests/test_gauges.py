import numpy as np

from foremast_1_b200 import _abi as A
from foremast_1_b200 import gauges


def test_exposition_names_and_values():
    rec = np.zeros(1, A.SERIES_REC)[0]
    rec["upper"], rec["lower"] = 12.5, 0.0
    text = gauges.exposition([("namespace_app_per_pod:http_server_requests_error_5xx", rec,
                               {"app": "demo", "namespace": "foremast-examples"}, [1700000060, 1700000120])])
    # names plotted by the UI: foremast-browser/src/config/metrics.js:14-24
    assert "# TYPE foremastbrain:namespace_app_per_pod:http_server_requests_error_5xx_upper gauge" in text
    assert 'foremastbrain:namespace_app_per_pod:http_server_requests_error_5xx_upper{app="demo",namespace="foremast-examples"} 12.5' in text
    assert 'foremastbrain:namespace_app_per_pod:http_server_requests_error_5xx_lower{app="demo",namespace="foremast-examples"} 0' in text
    # the anomaly gauge's value is the timestamp of the point (App.js:231-241)
    assert text.count("_anomaly{") == 2 and text.rstrip().endswith("1700000120")
