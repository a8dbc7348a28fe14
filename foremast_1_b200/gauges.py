"""Prometheus gauges the brain exposes on :8000/metrics.

The UI plots `foremastbrain:<metric>_upper`, `_lower` and `_anomaly`
(foremast-browser/src/config/metrics.js:7-28); the anomaly gauge's value is the
unix timestamp of the anomalous point (foremast-browser/src/App.js:231-241).
The numbers come straight out of the scoring records.

  gauge_lines / exposition   the text format
  GaugeRegistry              what the job loop feeds (brain.BrainWorker.poll_once): the latest bounds
                             and anomaly points of every metric of every job that is being watched
  MetricsServer              GET /metrics on port 8000, the endpoint the brain's ServiceMonitor
                             scrapes (deploy/foremast/3_brain/foremast-brain.yaml:87-122:
                             containerPort 8000, path /metrics)
"""
import http.server
import threading

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


class GaugeRegistry:
    """Latest gauges per job.  update() replaces a job's rows on every poll (a job is re-judged until
    its endTime); a finished job stays visible for `linger` further updates of the registry so that
    a scrape interval (30 s in the ServiceMonitor) cannot miss its last anomaly."""

    def __init__(self, linger=64):
        self._lock = threading.Lock()
        self._jobs = {}            # job id -> [rows, remaining updates or None]
        self._linger = linger

    def update(self, job_id, rows, final=False):
        with self._lock:
            for k in [k for k, v in self._jobs.items() if v[1] is not None]:
                self._jobs[k][1] -= 1
                if self._jobs[k][1] <= 0:
                    del self._jobs[k]
            self._jobs[job_id] = [list(rows), self._linger if final else None]

    def text(self):
        with self._lock:
            rows = [r for v in self._jobs.values() for r in v[0]]
        return exposition(rows) if rows else ""


class MetricsServer:
    """Serves registry.text() at /metrics (anything else: 404).  port=0 picks a free port (tests)."""

    def __init__(self, registry, host="0.0.0.0", port=8000):
        reg = registry

        class Handler(http.server.BaseHTTPRequestHandler):
            def do_GET(self):
                if self.path.split("?")[0] != "/metrics":
                    self.send_error(404)
                    return
                body = reg.text().encode()
                self.send_response(200)
                self.send_header("Content-Type", "text/plain; version=0.0.4; charset=utf-8")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def log_message(self, *a):
                pass

        self.httpd = http.server.ThreadingHTTPServer((host, port), Handler)
        self.port = self.httpd.server_address[1]
        self._thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)

    def start(self):
        self._thread.start()
        return self

    def stop(self):
        self.httpd.shutdown()
        self.httpd.server_close()
