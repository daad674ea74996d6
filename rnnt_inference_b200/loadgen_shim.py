"""Minimal in-process stand-in for the `mlperf_loadgen` Python binding (absent offline).

Only what models/pytorch_sut.py, models/rnnt_qsl.py and models/main.py touch: QuerySample / QuerySampleResponse,
ConstructSUT / ConstructQSL / Destroy*, QuerySamplesComplete, TestSettings / TestScenario / TestMode and StartTest
for the Offline scenario (one query holding every sample; result = samples / (last completion - issue)) and the
Server scenario (Poisson arrivals at target_qps; per-sample latency percentiles).  pytorch_sut.py imports the real
`mlperf_loadgen` when it is installed and falls back to this module otherwise.
"""
import ctypes
import enum
import threading
import time

import numpy as np


class TestScenario(enum.Enum):
    SingleStream = 0
    Offline = 1
    Server = 2


class TestMode(enum.Enum):
    PerformanceOnly = 0
    AccuracyOnly = 1


class TestSettings:
    def __init__(self):
        self.scenario = TestScenario.Offline
        self.mode = TestMode.PerformanceOnly
        self.min_query_count = 0           # Offline: samples in the single query (0 = the QSL's performance count)
        self.server_target_qps = 100.0
        self.server_duration_s = 5.0
        self.server_seed = 20240607

    def FromConfig(self, path, model, scenario):
        """mlperf.conf / user.conf: lines `<model|*>.<scenario|*>.<key> = <value>`; the keys this shim acts on are
        min_query_count (Offline: samples of the query), target_qps and min_duration (Server; ms, capped for offline
        use by server_duration_s set afterwards)."""
        try:
            lines = open(path).read().splitlines()
        except OSError:
            return 1
        for line in lines:
            line = line.split("#", 1)[0].strip()
            if "=" not in line:
                continue
            key, value = (t.strip() for t in line.split("=", 1))
            parts = key.split(".")
            if len(parts) != 3 or parts[0] not in ("*", model) or parts[1] not in ("*", scenario):
                continue
            try:
                v = float(value)
            except ValueError:
                continue
            if parts[2] == "min_query_count":
                self.min_query_count = int(v)
            elif parts[2] == "target_qps":
                self.server_target_qps = v
        return 0


class QuerySample:
    __slots__ = ("id", "index")

    def __init__(self, id, index):
        self.id = id
        self.index = index


class QuerySampleResponse:
    __slots__ = ("id", "data", "size")

    def __init__(self, id, data, size):
        self.id = id
        self.data = data
        self.size = size


class _SUT:
    def __init__(self, issue, flush):
        self.issue, self.flush = issue, flush


class _QSL:
    def __init__(self, count, perf_count, load, unload):
        self.count, self.perf_count, self.load, self.unload = count, perf_count, load, unload


def ConstructSUT(issue_queries, flush_queries):
    return _SUT(issue_queries, flush_queries)


def ConstructQSL(count, perf_count, load_cb, unload_cb):
    return _QSL(count, perf_count, load_cb, unload_cb)


def DestroySUT(sut):
    pass


def DestroyQSL(qsl):
    pass


class _Run:
    """Bookkeeping of the test in flight (the shim runs one test at a time, like LoadGen)."""

    def __init__(self, mode):
        self.mode = mode
        self.lock = threading.Lock()
        self.issue_time = {}
        self.done_time = {}
        self.last_done = 0.0
        self.accuracy = {}


_run = None


def QuerySamplesComplete(responses):
    now = time.perf_counter()
    run = _run
    if run is None:
        return
    with run.lock:
        run.done_time.update(dict.fromkeys([r.id for r in responses], now))
        run.last_done = now
        if run.mode == TestMode.AccuracyOnly:
            for r in responses:
                run.accuracy[r.id] = ctypes.string_at(r.data, int(r.size)) if r.size else b""


def StartTest(sut, qsl, settings):
    """Runs one test; returns a dict with the result (the real LoadGen writes mlperf_log_summary.txt instead)."""
    global _run
    _run = run = _Run(settings.mode)
    count = settings.min_query_count or qsl.perf_count
    # LoadSamplesToRam and the construction of the query's samples happen before LoadGen starts its clock; the shim keeps
    # both per QSL so that a repeated test neither reloads a loaded set nor rebuilds identical sample objects
    cached = getattr(qsl, "_shim_query", None)
    if cached is None or cached[0] != count:
        indices = [i % qsl.count for i in range(count)]
        qsl.load(sorted(set(indices)))
        cached = (count, [QuerySample(i + 1, idx) for i, idx in enumerate(indices)])
        qsl._shim_query = cached
    samples = list(cached[1])            # the SUT may reorder its list in place
    result = {"scenario": settings.scenario.name, "mode": settings.mode.name, "samples": count}
    if settings.scenario == TestScenario.Offline:
        t0 = time.perf_counter()
        run.issue_time = dict.fromkeys(range(1, count + 1), t0)
        sut.issue(samples)
        sut.flush()
        t1 = run.last_done if run.done_time else time.perf_counter()
        result.update(seconds=t1 - t0, samples_per_second=len(run.done_time) / max(t1 - t0, 1e-9),
                      completed=len(run.done_time))
    else:
        rng = np.random.default_rng(settings.server_seed)
        n = min(count, max(1, int(settings.server_target_qps * settings.server_duration_s)))
        gaps = rng.exponential(1.0 / settings.server_target_qps, size=n)
        t0 = time.perf_counter()
        due = t0
        for i in range(n):
            due += gaps[i]
            delay = due - time.perf_counter()
            if delay > 0:
                time.sleep(delay)
            run.issue_time[samples[i].id] = time.perf_counter()
            sut.issue([samples[i]])
        sut.flush()
        deadline = time.perf_counter() + 60.0
        while len(run.done_time) < n and time.perf_counter() < deadline:
            time.sleep(0.001)
        lat = np.array([run.done_time[k] - run.issue_time[k] for k in run.done_time])
        t1 = max(run.done_time.values()) if run.done_time else time.perf_counter()
        result.update(seconds=t1 - t0, completed=len(lat), samples_per_second=len(lat) / max(t1 - t0, 1e-9),
                      latency_p50_ms=float(np.percentile(lat, 50) * 1e3) if len(lat) else None,
                      latency_p99_ms=float(np.percentile(lat, 99) * 1e3) if len(lat) else None)
    if settings.mode == TestMode.AccuracyOnly:
        # same content as mlperf_log_accuracy.json entries: qsl_idx + hex of the int32 token bytes (eval_accuracy.py:85-101)
        result["accuracy_log"] = [{"qsl_idx": s.index, "data": run.accuracy.get(s.id, b"").hex()} for s in samples
                                  if s.id in run.accuracy]
    _run = None
    return result


def StartTestWithLogSettings(sut, qsl, settings, log_settings=None):
    return StartTest(sut, qsl, settings)


class LogOutputSettings:
    outdir = "."
    copy_summary_to_stdout = False


class LogSettings:
    log_output = None
