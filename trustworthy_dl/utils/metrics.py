"""``MetricsCollector`` — used by the reference (distributed_trainer.py:23,83,417,520;
experiment_runner.py:25,59) through ``collect_batch_metrics(dict)`` and ``get_summary()`` but not
shipped.  Metric families follow README.md:158-165: loss / perplexity / accuracy, detection rate
and false positives, trust evolution, throughput / latency / utilisation."""
from __future__ import annotations

import math
import time
from collections import defaultdict
from typing import Any, Dict, List, Optional

import numpy as np

__all__ = ["MetricsCollector"]


class MetricsCollector:
    def __init__(self, max_records: int = 100_000):
        self.max_records = int(max_records)
        self.records: List[Dict[str, Any]] = []
        self.trust_evolution: Dict[int, List[float]] = defaultdict(list)
        self.events: List[Dict[str, Any]] = []
        self._t0 = time.time()
        self._step_times: List[float] = []
        self._tokens = 0.0

    def collect_batch_metrics(self, metrics: Dict[str, Any]) -> None:
        rec = dict(metrics)
        rec.setdefault("timestamp", time.time())
        ts = rec.get("trust_scores")
        if ts:
            for node, val in ts.items():
                self.trust_evolution[int(node)].append(float(val))
        if "step_time" in rec:
            self._step_times.append(float(rec["step_time"]))
        self._tokens += float(rec.get("samples", 0.0))
        if len(self.records) < self.max_records:
            self.records.append(rec)

    def record_event(self, kind: str, **info) -> None:
        self.events.append({"kind": kind, "timestamp": time.time(), **info})

    def record_detection(self, flagged: bool, ground_truth: Optional[bool]) -> None:
        self.record_event("detection", flagged=bool(flagged), truth=ground_truth)

    def losses(self) -> List[float]:
        return [float(r["loss"]) for r in self.records if "loss" in r and r["loss"] is not None]

    def get_summary(self) -> Dict[str, Any]:
        losses = self.losses()
        out: Dict[str, Any] = {"num_batches": len(self.records), "wall_time": time.time() - self._t0}
        if losses:
            arr = np.asarray(losses, dtype=np.float64)
            fin = arr[np.isfinite(arr)]                 # a tainted step can report a non-finite loss (its update is dropped)
            out["non_finite_losses"] = int(arr.size - fin.size)
            if fin.size:
                out.update({"mean_loss": float(fin.mean()), "final_loss": float(fin[-1]), "min_loss": float(fin.min()),
                            "perplexity": float(math.exp(min(20.0, fin[-min(10, fin.size):].mean())))})
        accs = [r["accuracy"] for r in self.records if "accuracy" in r]
        if accs:
            out["accuracy"] = float(np.mean(accs[-10:]))
        if self._step_times:
            st = np.asarray(self._step_times)
            out.update({"mean_step_time": float(st.mean()), "p50_step_time": float(np.median(st)),
                        "p99_step_time": float(np.percentile(st, 99))})
            if self._tokens:
                out["throughput"] = float(self._tokens / st.sum())
        det = [e for e in self.events if e["kind"] == "detection" and e.get("truth") is not None]
        if det:
            tp = sum(e["flagged"] and e["truth"] for e in det)
            fp = sum(e["flagged"] and not e["truth"] for e in det)
            pos = sum(bool(e["truth"]) for e in det)
            neg = len(det) - pos
            out["detection_rate"] = tp / pos if pos else None
            out["false_positive_rate"] = fp / neg if neg else None
        if self.trust_evolution:
            out["final_trust"] = {k: v[-1] for k, v in self.trust_evolution.items()}
            out["min_trust"] = {k: float(min(v)) for k, v in self.trust_evolution.items()}
        out["events"] = {k: sum(1 for e in self.events if e["kind"] == k) for k in {e["kind"] for e in self.events}}
        return out

    def reset(self) -> None:
        self.records.clear()
        self.trust_evolution.clear()
        self.events.clear()
        self._step_times.clear()
        self._tokens = 0.0
