"""Prometheus gauges the brain exposes on :8000/metrics.

The UI plots `foremastbrain:<metric>_upper`, `_lower` and `_anomaly`
(foremast-browser/src/config/metrics.js:7-28); the anomaly gauge's value is the
unix timestamp of the anomalous point (foremast-browser/src/App.js:231-241).
The numbers come straight out of the scoring records; this module only formats
the text exposition.  Serving it is one line with prometheus_client or
http.server in the embedding process (deploy/foremast/3_brain/foremast-brain.yaml:87-122).
"""

PREFIX = "foremastbrain:"


def _labels(labels):
    if not labels:
        return ""
    return "{" + ",".join(f'{k}="{str(v)}"' for k, v in sorted(labels.items())) + "}"


def gauge_lines(metric_name, rec, labels=None, anomaly_timestamps=()):
    """Exposition lines for one scored series.  `metric_name` is the recording-rule
    name (e.g. namespace_app_per_pod:http_server_requests_error_5xx)."""
    lab = _labels(labels)
    base = PREFIX + metric_name
    out = [f"{base}_upper{lab} {float(rec['upper']):.10g}",
           f"{base}_lower{lab} {float(rec['lower']):.10g}"]
    for ts in anomaly_timestamps:
        out.append(f"{base}_anomaly{lab} {int(ts)}")
    return out


def exposition(scored):
    """scored: iterable of (metric_name, record, labels, anomaly_timestamps)."""
    names, lines = set(), []
    for name, rec, labels, ts in scored:
        for suffix in ("_upper", "_lower", "_anomaly"):
            full = PREFIX + name + suffix
            if full not in names:
                names.add(full)
                lines.append(f"# TYPE {full} gauge")
        lines += gauge_lines(name, rec, labels, ts)
    return "\n".join(lines) + "\n"
