"""Python mirror of the reference's pipeline stage for the decode path.

The reference's stage is Java: ``TFModel`` (src/main/java/org/apache/flink/table/ml/lib/tensorflow/
TFModel.java:29-87) with the param mix-ins ``HasClusterConfig``, ``HasInferencePythonConfig``,
``HasInferenceSelectedCols/OutputCols/OutputTypes`` (param/*.java).  No JDK exists in this image, so the
stage-level contract is mirrored here with the same names, defaults, JSON round trip and error behaviour,
on top of the C ABI (``native.Decoder``).  ``transform`` takes an iterable of rows (dicts) instead of a
Flink ``Table``; everything below the stage -- the operator body -- is ``libpgen_cuda.so``.

Inert params (accepted, stored, ignored -- there is no Python/TensorFlow worker to start):
zookeeper_connect_str, ps_num, inference_scripts, inference_map_func, inference_env_path.
"""
from __future__ import annotations

import json
from typing import Dict, Iterable, Iterator, List, Optional, Sequence

from . import native, sharding

STRING = "STRING"            # DataTypes.STRING (util/CodingUtils.java:37-62)

_PARAM_DEFAULTS = {          # name -> (required, has_default, default)   param/*.java
    "zookeeper_connect_str": (True, True, "127.0.0.1:2181"),       # HasClusterConfig.java:15-19
    "worker_num": (True, True, 1),                                 # :20-24
    "ps_num": (True, True, 0),                                     # :25-29
    "inference_scripts": (True, False, None),                      # HasInferencePythonConfig.java:17-20
    "inference_map_func": (True, False, None),                     # :21-24
    "inference_hyper_params_key": (True, False, None),             # :25-29
    "inference_hyper_params": (True, True, []),                    # :30-34
    "inference_env_path": (False, True, None),                     # :35-39
    "inference_selected_cols": (True, False, None),
    "inference_output_cols": (True, False, None),
    "inference_output_types": (True, False, None),
}


class TFModel:
    """``new TFModel().setWorkerNum(1)....setInferenceHyperParams(...).transform(rows)``
    (usage: src/test/java/.../TensorFlowTest.java:219-234, 93-107)."""

    def __init__(self):
        self._params: Dict[str, object] = {}

    # ---- Params plumbing (flink-ml-api WithParams) ----
    def _get(self, name: str):
        if name in self._params:
            return self._params[name]
        required, has_default, default = _PARAM_DEFAULTS[name]
        if has_default:
            return list(default) if isinstance(default, list) else default
        raise RuntimeError("Cannot find default value for param '%s'" % name)

    def _set(self, name: str, value) -> "TFModel":
        self._params[name] = value
        return self

    def getParams(self) -> Dict[str, object]:                      # TFModel.java:83-86
        return self._params

    def toJson(self) -> str:                                       # TensorFlowTest.java:152
        return json.dumps(self._params, sort_keys=True)

    def loadJson(self, s: str) -> "TFModel":                      # TensorFlowTest.java:162-163
        d = json.loads(s)
        unknown = set(d) - set(_PARAM_DEFAULTS)
        if unknown:
            raise RuntimeError("unknown params in JSON: %s" % sorted(unknown))
        self._params = d
        return self

    # ---- HasClusterConfig ----
    def getZookeeperConnStr(self): return self._get("zookeeper_connect_str")
    def setZookeeperConnStr(self, v: str): return self._set("zookeeper_connect_str", v)
    def getWorkerNum(self) -> int: return int(self._get("worker_num"))
    def setWorkerNum(self, v: int): return self._set("worker_num", int(v))
    def getPsNum(self) -> int: return int(self._get("ps_num"))
    def setPsNum(self, v: int): return self._set("ps_num", int(v))

    # ---- HasInferencePythonConfig ----
    def getInferenceScripts(self): return self._get("inference_scripts")
    def setInferenceScripts(self, v: Sequence[str]): return self._set("inference_scripts", list(v))
    def getInferenceMapFunc(self): return self._get("inference_map_func")
    def setInferenceMapFunc(self, v: str): return self._set("inference_map_func", v)
    def getInferenceHyperParamsKey(self): return self._get("inference_hyper_params_key")
    def setInferenceHyperParamsKey(self, v: str): return self._set("inference_hyper_params_key", v)
    def getInferenceHyperParams(self) -> List[str]: return self._get("inference_hyper_params")
    def setInferenceHyperParams(self, v: Sequence[str]): return self._set("inference_hyper_params", list(v))
    def getInferenceEnvPath(self): return self._get("inference_env_path")
    def setInferenceEnvPath(self, v: Optional[str]): return self._set("inference_env_path", v)

    # ---- HasInferenceSelectedCols / OutputCols / OutputTypes ----
    def getInferenceSelectedCols(self): return self._get("inference_selected_cols")
    def setInferenceSelectedCols(self, v: Sequence[str]): return self._set("inference_selected_cols", list(v))
    def getInferenceOutputCols(self): return self._get("inference_output_cols")
    def setInferenceOutputCols(self, v: Sequence[str]): return self._set("inference_output_cols", list(v))
    def getInferenceOutputTypes(self): return self._get("inference_output_types")
    def setInferenceOutputTypes(self, v: Sequence[str]): return self._set("inference_output_types", list(v))

    # ---- the stage ----
    def hyperparams_string(self) -> str:
        """String.join(" ", getInferenceHyperParams()) -- what TFModel puts under the hyper-params key
        (TFModel.java:47) and main_on_flink splits again (run_summarization.py:418-420)."""
        return " ".join(self.getInferenceHyperParams())

    def _check_schema(self):
        sel, out, types = self.getInferenceSelectedCols(), self.getInferenceOutputCols(), self.getInferenceOutputTypes()
        for need in ("uuid", "article", "reference"):              # features of batcher.py:554-557
            if need not in sel:
                raise RuntimeError("inferenceSelectedCols must contain '%s' (got %s)" % (need, sel))
        if len(out) != len(types):
            raise RuntimeError("inferenceOutputCols and inferenceOutputTypes differ in length")
        for c, t in zip(out, types):
            if c not in ("uuid", "article", "summary", "reference"):   # flink_writer.py:26-33
                raise RuntimeError("unknown output column '%s'" % c)
            if t != STRING:
                raise RuntimeError("output column '%s' must be STRING" % c)
        return sel, out

    def transform(self, rows: Iterable[dict], devices: Optional[Sequence[int]] = None) -> Iterator[dict]:
        """TFModel.transform (TFModel.java:55-81): select the input columns (:34-36), run the operator at
        parallelism worker_num (one handle = one subtask = one GPU), emit rows with the output schema
        (:38-41) in input order.  Any failure surfaces as RuntimeError, like the reference's
        RuntimeException wrapper (:78-80)."""
        try:
            sel, out_cols = self._check_schema()
            n_workers = self.getWorkerNum()
            if n_workers < 1:
                raise RuntimeError("worker_num must be >= 1")
            hyper = self.hyperparams_string()
            if devices is None:
                devices = list(range(n_workers))
            workers = [native.Decoder(hyper, device=devices[i % len(devices)]) for i in range(n_workers)]
        except native.PgenError as e:
            raise RuntimeError(str(e)) from e
        try:
            # rows are dealt round-robin to the subtasks; each subtask returns its rows in submission order, so
            # popping the subtasks' queues in the same round-robin order restores the input order.
            queues: List[List[tuple]] = [[] for _ in workers]
            submitted, emitted = 0, 0

            def drain(flush: bool):
                nonlocal emitted
                for w, q in zip(workers, queues):
                    q.extend(w.poll(flush=flush))
                while queues[sharding.owner(emitted, n_workers)]:
                    uuid, article, summary, reference = queues[sharding.owner(emitted, n_workers)].pop(0)
                    full = {"uuid": uuid, "article": article, "summary": summary, "reference": reference}
                    emitted += 1
                    yield {c: full[c] for c in out_cols}

            for row in rows:
                vals = {}
                for c in sel:
                    if c not in row:
                        raise RuntimeError("input row lacks selected column '%s'" % c)
                    v = row[c]
                    vals[c] = v.encode() if isinstance(v, str) else bytes(v)
                workers[sharding.owner(submitted, n_workers)].submit(vals["uuid"], vals["article"], vals["reference"])
                submitted += 1
                yield from drain(False)
            yield from drain(True)
            assert emitted == submitted
        except native.PgenError as e:
            raise RuntimeError(str(e)) from e
        finally:
            for w in workers:
                w.close()
