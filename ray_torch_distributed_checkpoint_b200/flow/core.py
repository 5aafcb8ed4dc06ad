"""Flow runner core: ``FlowSpec``, ``@step``, ``Parameter``, ``current``, ``Run``/``Task``, decorators, CLI.

Datastore layout (``$RTDC_FLOW_DATASTORE`` or ``~/.rtdc_flow``)::

    <root>/<FlowName>/<run_id>/_meta.json                 status, parameters, trigger info
    <root>/<FlowName>/<run_id>/<step>/<task_id>/artifacts.pkl   cloudpickled ``self.*`` attributes
    <root>/<FlowName>/<run_id>/<step>/<task_id>/ray_storage/    ``current.ray_storage_path`` of that task attempt
    <root>/<FlowName>/<run_id>/<step>/<task_id>/cards/<id>.html
    <root>/_deployments/<FlowName>.json                   written by ``argo-workflows create``

``Run("Flow/run_id").data.result`` and ``Task("Flow/run_id/step/task_id").data.result`` resolve pickled
artifacts exactly as the reference's flows expect (``/root/reference/train_flow.py:68-73``,
``/root/reference/eval_flow.py:40-54``).
"""
from __future__ import annotations

import argparse
import functools
import importlib.util
import json
import os
import sys
import time
import traceback
from typing import Any, Callable, Dict, List, Optional

import cloudpickle

from .cards import CardCollection


class FlowError(RuntimeError):
    pass


def datastore_root() -> str:
    return os.path.abspath(os.path.expanduser(os.environ.get("RTDC_FLOW_DATASTORE", "~/.rtdc_flow")))


_namespace = [None]


def namespace(ns: Optional[str]):
    """Kept for API parity (single-user datastore: namespaces do not partition anything here)."""
    _namespace[0] = ns
    return ns


# ------------------------------------------------------------------------------------------------ current
class _Trigger:
    def __init__(self, run: "Run"):
        self.run = run
        self.runs = [run]


class _Parallel:
    def __init__(self, node_index=0, num_nodes=1):
        self.node_index, self.num_nodes = node_index, num_nodes
        self.main_ip = "127.0.0.1"


class _Current:
    """Per-task context, the ``metaflow.current`` singleton."""

    def __init__(self):
        self._reset()

    def _reset(self):
        self.flow_name = None
        self.run_id = None
        self.step_name = None
        self.task_id = None
        self.retry_count = 0
        self.origin_run_id = None
        self.parallel = _Parallel()
        self.card = CardCollection()
        self._task_dir = None
        self._trigger = None
        self._has_ray = False

    @property
    def pathspec(self) -> str:
        return f"{self.flow_name}/{self.run_id}/{self.step_name}/{self.task_id}"

    @property
    def ray_storage_path(self) -> str:
        """Unique per task attempt, inside the datastore (ref: README.md:13-15). Available in steps
        decorated with ``@metaflow_ray``."""
        if not self._has_ray:
            raise AttributeError("current.ray_storage_path is only available inside a @metaflow_ray step")
        p = os.path.join(self._task_dir, "ray_storage", f"attempt_{self.retry_count}")
        os.makedirs(p, exist_ok=True)
        return p

    @property
    def trigger(self):
        if self._trigger is None:
            raise AttributeError("this run was not triggered by another flow")
        return self._trigger

    @property
    def is_running_flow(self) -> bool:
        return self.flow_name is not None


current = _Current()


# ------------------------------------------------------------------------------------------------ parameters / decorators
class Parameter:
    def __init__(self, name: str, default: Any = None, help: str = "", type: Optional[type] = None, required: bool = False):
        self.name, self.default, self.help, self.required = name, default, help, required
        self.type = type or (builtins_type(default) if default is not None else str)
        self.attr = None

    def __set_name__(self, owner, attr):
        self.attr = attr

    def __get__(self, obj, objtype=None):
        if obj is None:
            return self
        return obj.__dict__.get("_param_" + self.attr, self.default)


def builtins_type(v):
    if isinstance(v, bool):
        return lambda s: s if isinstance(s, bool) else str(s).lower() in ("1", "true", "yes")
    return type(v)


def step(fn: Callable) -> Callable:
    fn._is_step = True
    fn._decos = getattr(fn, "_decos", {})
    return fn


def _deco(name: str, **defaults):
    """Decorator factory usable as ``@name`` and ``@name(k=v)``; stores its options on the function."""
    def outer(*args, **kwargs):
        def apply(fn):
            d = dict(getattr(fn, "_decos", {}))
            opts = dict(defaults)
            opts.update(kwargs)
            d[name] = opts
            fn._decos = d
            return fn
        if len(args) == 1 and callable(args[0]) and not kwargs:
            return apply(args[0])
        return apply
    return outer


retry = _deco("retry", times=3, minutes_between_retries=0)
metaflow_ray = _deco("metaflow_ray", all_nodes_started_timeout=300)
kubernetes = _deco("kubernetes")
pypi = _deco("pypi")
card = _deco("card", type="blank", id="default")


def gpu_profile(*args, **kwargs):
    """``@gpu_profile(interval=1)``: sample GPU utilisation / memory in a background thread while the step
    runs and attach the series as a card (ref: train_flow.py:51, eval_flow.py:57)."""
    return _deco("gpu_profile", interval=1)(*args, **kwargs)


def _class_deco(name: str, **defaults):
    def outer(*args, **kwargs):
        def apply(cls):
            d = dict(getattr(cls, "_flow_decos", {}))
            opts = dict(defaults)
            opts.update(kwargs)
            d[name] = opts
            cls._flow_decos = d
            return cls
        if len(args) == 1 and isinstance(args[0], type) and not kwargs:
            return apply(args[0])
        return apply
    return outer


schedule = _class_deco("schedule", cron=None)
trigger_on_finish = _class_deco("trigger_on_finish", flow=None)


# ------------------------------------------------------------------------------------------------ client API
class _Data:
    def __init__(self, artifacts: Dict[str, Any]):
        self.__dict__["_a"] = artifacts

    def __getattr__(self, k):
        try:
            return self.__dict__["_a"][k]
        except KeyError:
            raise AttributeError(k) from None

    def __contains__(self, k):
        return k in self._a


def _load_artifacts(task_dir: str) -> Dict[str, Any]:
    p = os.path.join(task_dir, "artifacts.pkl")
    if not os.path.exists(p):
        return {}
    with open(p, "rb") as f:
        return cloudpickle.load(f)


class Task:
    def __init__(self, pathspec: str):
        parts = pathspec.strip("/").split("/")
        if len(parts) != 4:
            raise FlowError(f"task pathspec must look like Flow/run_id/step/task_id (got {pathspec!r})")
        self.pathspec = "/".join(parts)
        self.flow_name, self.run_id, self.step_name, self.id = parts
        self._dir = os.path.join(datastore_root(), *parts)
        if not os.path.isdir(self._dir):
            raise FlowError(f"no such task in the datastore: {self.pathspec}")

    @property
    def data(self) -> _Data:
        return _Data(_load_artifacts(self._dir))

    @property
    def successful(self) -> bool:
        return os.path.exists(os.path.join(self._dir, "_ok"))

    def __repr__(self):
        return f"Task('{self.pathspec}')"


class Step:
    def __init__(self, pathspec: str):
        parts = pathspec.strip("/").split("/")
        self.pathspec = "/".join(parts)
        self._dir = os.path.join(datastore_root(), *parts)
        if not os.path.isdir(self._dir):
            raise FlowError(f"no such step in the datastore: {self.pathspec}")

    def tasks(self) -> List[Task]:
        return [Task(f"{self.pathspec}/{t}") for t in sorted(os.listdir(self._dir), key=lambda s: int(s) if s.isdigit() else 0)]

    @property
    def task(self) -> Task:
        return self.tasks()[0]


class Run:
    def __init__(self, pathspec: str):
        parts = pathspec.strip("/").split("/")
        if len(parts) != 2:
            raise FlowError(f"run pathspec must look like Flow/run_id (got {pathspec!r})")
        self.pathspec = "/".join(parts)
        self.flow_name, self.id = parts
        self._dir = os.path.join(datastore_root(), *parts)
        if not os.path.isdir(self._dir):
            raise FlowError(f"no such run in the datastore: {self.pathspec}")

    def _meta(self) -> Dict[str, Any]:
        with open(os.path.join(self._dir, "_meta.json")) as f:
            return json.load(f)

    @property
    def successful(self) -> bool:
        return self._meta().get("status") == "success"

    @property
    def finished(self) -> bool:
        return self._meta().get("status") in ("success", "failed")

    def steps(self) -> List[Step]:
        order = self._meta().get("step_order", [])
        return [Step(f"{self.pathspec}/{s}") for s in order if os.path.isdir(os.path.join(self._dir, s))]

    def __getitem__(self, step_name: str) -> Step:
        return Step(f"{self.pathspec}/{step_name}")

    @property
    def end_task(self) -> Task:
        return self["end"].task

    @property
    def data(self) -> _Data:
        """Artifacts of the run's ``end`` task (what ``Run(...).data.result`` reads, train_flow.py:73)."""
        return self.end_task.data

    def __repr__(self):
        return f"Run('{self.pathspec}')"


class Flow:
    def __init__(self, name: str):
        self.name = name
        self._dir = os.path.join(datastore_root(), name)

    def runs(self) -> List[Run]:
        if not os.path.isdir(self._dir):
            return []
        ids = sorted((d for d in os.listdir(self._dir) if d.isdigit()), key=int, reverse=True)
        return [Run(f"{self.name}/{i}") for i in ids]

    @property
    def latest_run(self) -> Optional[Run]:
        r = self.runs()
        return r[0] if r else None

    @property
    def latest_successful_run(self) -> Optional[Run]:
        for r in self.runs():
            if r.successful:
                return r
        return None


# ------------------------------------------------------------------------------------------------ GPU sampler
class _GpuSampler:
    def __init__(self, interval: float):
        import threading
        self.interval = max(0.05, float(interval))
        self.samples: List[Dict[str, Any]] = []
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            handles = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in range(pynvml.nvmlDeviceGetCount())]
        except Exception:
            return
        t0 = time.time()
        while not self._stop.is_set():
            row = {"t": round(time.time() - t0, 3)}
            for i, h in enumerate(handles):
                try:
                    u = pynvml.nvmlDeviceGetUtilizationRates(h)
                    m = pynvml.nvmlDeviceGetMemoryInfo(h)
                    row[f"gpu{i}_util"] = u.gpu
                    row[f"gpu{i}_mem_mib"] = m.used // (1 << 20)
                    row[f"gpu{i}_sm_mhz"] = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                    row[f"gpu{i}_power_w"] = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pass
            self.samples.append(row)
            self._stop.wait(self.interval)

    def start(self):
        self._thread.start()

    def stop(self):
        self._stop.set()
        self._thread.join(timeout=2.0)


def _gpu_profile_card(samples: List[Dict[str, Any]]):
    """Attach the sampler's series to the task as the ``gpu_profile`` card (peak / mean per GPU)."""
    from .cards import Markdown, Table
    c = current.card["gpu_profile"]
    c.append(Markdown(f"### GPU profile ({len(samples)} samples)"))
    if not samples:
        c.append(Markdown("no NVML device visible to this task"))
        return
    keys = sorted(k for k in samples[-1] if k != "t")
    rows = []
    for k in keys:
        vals = [s[k] for s in samples if k in s]
        rows.append([k, f"{sum(vals) / len(vals):.1f}", f"{max(vals):.1f}"])
    c.append(Table(headers=["series", "mean", "max"], data=rows))


# ------------------------------------------------------------------------------------------------ FlowSpec
class _Inputs(list):
    """``inputs`` of a join step: iterable of per-task namespaces; attribute access picks by step name."""


class _InputView:
    def __init__(self, artifacts: Dict[str, Any]):
        self.__dict__.update(artifacts)


class FlowSpec:
    """Subclass, decorate methods with ``@step``, call ``MyFlow()`` under ``if __name__ == '__main__'``."""

    _flow_decos: Dict[str, Any] = {}

    def __init__(self, use_cli: bool = True, argv: Optional[List[str]] = None):
        self._transition = None
        self._artifacts_snapshot = None
        if use_cli:
            self._main(sys.argv[1:] if argv is None else argv)

    # -------------------------------------------------------------- graph helpers
    @classmethod
    def _parameters(cls) -> Dict[str, Parameter]:
        out = {}
        for klass in reversed(cls.__mro__):
            for k, v in vars(klass).items():
                if isinstance(v, Parameter):
                    out[k] = v
        return out

    def next(self, *targets, num_parallel: Optional[int] = None, foreach: Optional[str] = None):
        if foreach is not None:
            raise FlowError("foreach transitions are not supported by this runner")
        names = []
        for t in targets:
            if not getattr(t, "_is_step", False):
                raise FlowError("self.next() takes @step methods")
            names.append(t.__name__)
        self._transition = {"targets": names, "num_parallel": num_parallel}

    def _artifacts(self) -> Dict[str, Any]:
        skip = {"_transition", "_artifacts_snapshot"}
        out = {}
        for k, v in self.__dict__.items():
            if k in skip or k.startswith("_param_"):
                continue
            out[k] = v
        for attr, _p in self._parameters().items():
            out[attr] = getattr(self, attr)
        return out

    # -------------------------------------------------------------- execution
    @classmethod
    def _new_run_dir(cls) -> (str, str):
        root = os.path.join(datastore_root(), cls.__name__)
        os.makedirs(root, exist_ok=True)
        while True:
            run_id = str(int(time.time() * 1000000))
            d = os.path.join(root, run_id)
            try:
                os.makedirs(d)
                return run_id, d
            except FileExistsError:
                time.sleep(0.001)

    def _run_task(self, run_dir: str, step_name: str, task_id: int, inputs: Optional[_Inputs], parallel: _Parallel,
                  execute_body: bool = True) -> Dict[str, Any]:
        fn = getattr(type(self), step_name)
        decos = getattr(fn, "_decos", {})
        task_dir = os.path.join(run_dir, step_name, str(task_id))
        os.makedirs(task_dir, exist_ok=True)
        times = int(decos.get("retry", {}).get("times", 0)) if "retry" in decos else 0
        attempt = 0
        base_state = dict(self.__dict__)
        while True:
            self.__dict__.clear()
            self.__dict__.update(base_state)
            self._transition = None
            current.step_name, current.task_id, current.retry_count = step_name, str(task_id), attempt
            current.parallel = parallel
            current.card = CardCollection()
            current._task_dir = task_dir
            current._has_ray = "metaflow_ray" in decos
            sampler = None
            if "gpu_profile" in decos:
                sampler = _GpuSampler(decos["gpu_profile"].get("interval", 1))
                sampler.start()
            try:
                if execute_body:
                    if inputs is not None:
                        fn(self, inputs)
                    else:
                        fn(self)
                else:
                    self._transition = {"targets": [], "num_parallel": None, "worker": True}
                break
            except Exception as e:
                if attempt < times:
                    print(f"[flow] step {step_name} (task {task_id}) failed on attempt {attempt}: {e!r}; retrying "
                          f"({attempt + 1}/{times})", flush=True)
                    attempt += 1
                    continue
                with open(os.path.join(task_dir, "_error.txt"), "w") as f:
                    f.write(traceback.format_exc())
                raise
            finally:
                if sampler is not None:
                    sampler.stop()
                    with open(os.path.join(task_dir, "gpu_profile.json"), "w") as f:
                        json.dump(sampler.samples, f)
                    _gpu_profile_card(sampler.samples)
        arts = self._artifacts() if execute_body else {}
        with open(os.path.join(task_dir, "artifacts.pkl"), "wb") as f:
            cloudpickle.dump(arts, f)
        cards_dir = os.path.join(task_dir, "cards")
        for cid, c in current.card.items():
            os.makedirs(cards_dir, exist_ok=True)
            with open(os.path.join(cards_dir, f"{cid}.html"), "w") as f:
                f.write(c.render(title=f"{type(self).__name__}/{current.run_id}/{step_name} — {cid}"))
        open(os.path.join(task_dir, "_ok"), "w").close()
        return arts

    def _execute(self, params: Dict[str, Any], trigger_run: Optional[Run] = None) -> str:
        cls = type(self)
        if not getattr(getattr(cls, "start", None), "_is_step", False) or not getattr(getattr(cls, "end", None), "_is_step", False):
            raise FlowError("a flow needs @step methods named 'start' and 'end'")
        run_id, run_dir = cls._new_run_dir()
        for attr, p in cls._parameters().items():
            self.__dict__["_param_" + attr] = params.get(attr, p.default)
        current._reset()
        current.flow_name, current.run_id = cls.__name__, run_id
        current._trigger = _Trigger(trigger_run) if trigger_run is not None else None
        meta = {"flow": cls.__name__, "run_id": run_id, "status": "running", "started": time.time(),
                "parameters": {k: repr(v) for k, v in params.items()}, "step_order": [],
                "triggered_by": trigger_run.pathspec if trigger_run else None}

        def save_meta():
            with open(os.path.join(run_dir, "_meta.json"), "w") as f:
                json.dump(meta, f, indent=1)

        save_meta()
        print(f"[flow] {cls.__name__}/{run_id} started (datastore {datastore_root()})", flush=True)
        task_counter = [0]

        def new_task_id():
            task_counter[0] += 1
            return task_counter[0]

        try:
            step_name, pending_inputs = "start", None
            while True:
                meta["step_order"].append(step_name)
                save_meta()
                arts = self._run_task(run_dir, step_name, new_task_id(), pending_inputs, _Parallel())
                tr = self._transition
                if step_name == "end":
                    break
                if not tr or not tr["targets"]:
                    raise FlowError(f"step {step_name} did not call self.next()")
                if len(tr["targets"]) != 1:
                    raise FlowError("static branches are not supported by this runner (one target per self.next)")
                nxt = tr["targets"][0]
                pending_inputs = None
                if tr.get("num_parallel"):
                    n = int(tr["num_parallel"])
                    gang_fn = getattr(cls, nxt)
                    is_ray = "metaflow_ray" in getattr(gang_fn, "_decos", {})
                    meta["step_order"].append(nxt)
                    save_meta()
                    base = dict(self.__dict__)
                    gathered = []
                    for node in range(n):
                        self.__dict__.clear()
                        self.__dict__.update(base)
                        # @metaflow_ray: the tasks form one cluster and the step body runs on the head
                        # only (train_flow.py:42; explains the try/except in its join, :84-88)
                        a = self._run_task(run_dir, nxt, new_task_id(), None, _Parallel(node, n),
                                           execute_body=(node == 0) or not is_ray)
                        gathered.append(_InputView(a))
                        if node == 0:
                            head_tr = self._transition
                    self.__dict__.clear()
                    self.__dict__.update(base)
                    if not head_tr or not head_tr["targets"]:
                        raise FlowError(f"step {nxt} did not call self.next()")
                    step_name = head_tr["targets"][0]
                    pending_inputs = _Inputs(gathered)
                    # a join starts from a clean namespace (plus parameters), like Metaflow's
                    keep = {k: v for k, v in self.__dict__.items() if k.startswith("_")}
                    self.__dict__.clear()
                    self.__dict__.update(keep)
                    continue
                step_name = nxt
            meta["status"] = "success"
        except BaseException as e:
            meta["status"] = "failed"
            meta["error"] = repr(e)
            raise
        finally:
            meta["finished"] = time.time()
            save_meta()
            current._reset()
        print(f"[flow] {cls.__name__}/{run_id} finished successfully", flush=True)
        return run_id

    # -------------------------------------------------------------- CLI
    def _main(self, argv: List[str]):
        cls = type(self)
        ap = argparse.ArgumentParser(prog=sys.argv[0], description=cls.__doc__)
        ap.add_argument("--environment", default=None, help="accepted for command-line parity; ignored on one box")
        ap.add_argument("--datastore", default=None, help="datastore root (default $RTDC_FLOW_DATASTORE or ~/.rtdc_flow)")
        sub = ap.add_subparsers(dest="cmd")

        def add_params(sp):
            for attr, p in cls._parameters().items():
                flag = "--" + p.name
                alts = {flag, "--" + p.name.replace("_", "-"), "--" + p.name.replace("-", "_")}
                sp.add_argument(*sorted(alts), dest=attr, default=p.default, type=p.type if p.type is not bool else None,
                                help=p.help or None, required=p.required)

        for name in ("run", "evaluate"):
            add_params(sub.add_parser(name, help="execute the flow locally"))
        argo = sub.add_parser("argo-workflows", help="deployment commands (local emulation)")
        argo_sub = argo.add_subparsers(dest="argo_cmd")
        argo_sub.add_parser("create")
        add_params(argo_sub.add_parser("trigger"))
        serve = argo_sub.add_parser("serve", help="honour @schedule(cron=...): run the flow at every matching minute")
        add_params(serve)
        serve.add_argument("--max-runs", type=int, default=0, help="stop after this many scheduled runs (0: until interrupted)")
        serve.add_argument("--poll-seconds", type=float, default=1.0)
        serve.add_argument("--now", default=None, help="pretend the clock starts at this ISO time (tests)")
        sub.add_parser("show")
        args = ap.parse_args(argv)
        if args.datastore:
            os.environ["RTDC_FLOW_DATASTORE"] = args.datastore
        params = {attr: getattr(args, attr) for attr in cls._parameters() if hasattr(args, attr)}
        if args.cmd in ("run", "evaluate"):
            self._execute(params)
        elif args.cmd == "argo-workflows":
            if args.argo_cmd == "create":
                self._deploy()
            elif args.argo_cmd == "trigger":
                run_id = self._execute(params)
                self._fire_downstream(Run(f"{cls.__name__}/{run_id}"))
            elif args.argo_cmd == "serve":
                self._serve(params, args.max_runs, args.poll_seconds, args.now)
            else:
                ap.error("argo-workflows needs 'create', 'trigger' or 'serve'")
        elif args.cmd == "show":
            print(cls.__name__, "steps:", [k for k, v in vars(cls).items() if getattr(v, "_is_step", False)])
            print("parameters:", {p.name: p.default for p in cls._parameters().values()})
            print("decorators:", getattr(cls, "_flow_decos", {}))
        else:
            ap.print_help()

    # -------------------------------------------------------------- deployment emulation
    @classmethod
    def _deploy(cls):
        d = os.path.join(datastore_root(), "_deployments")
        os.makedirs(d, exist_ok=True)
        rec = {"flow": cls.__name__, "file": os.path.abspath(sys.modules[cls.__module__].__file__),
               "decorators": getattr(cls, "_flow_decos", {}), "created": time.time()}
        with open(os.path.join(d, f"{cls.__name__}.json"), "w") as f:
            json.dump(rec, f, indent=1)
        sched = rec["decorators"].get("schedule", {}).get("cron")
        trig = rec["decorators"].get("trigger_on_finish", {}).get("flow")
        print(f"[flow] deployed {cls.__name__}" + (f" (cron {sched!r})" if sched else "") + (f" (runs after {trig})" if trig else ""))

    def _serve(self, params: Dict[str, Any], max_runs: int = 0, poll_s: float = 1.0, now: Optional[str] = None):
        """The scheduler a deployment would get from Argo's CronWorkflow, on this box: at every minute that matches the
        flow's ``@schedule(cron=...)`` (ref train_flow.py:20, ``*/5 * * * *``) run the flow once, then the flows deployed with
        ``@trigger_on_finish`` on it. A run that outlasts its slot makes the scheduler skip the slots it overlapped."""
        import datetime as dt
        cls = type(self)
        expr = getattr(cls, "_flow_decos", {}).get("schedule", {}).get("cron")
        if not expr:
            raise FlowError(f"{cls.__name__} has no @schedule(cron=...)")
        cron = CronSchedule(expr)
        self._deploy()
        offset = (dt.datetime.fromisoformat(now) - dt.datetime.now()) if now else dt.timedelta(0)
        runs, last_slot = 0, None
        print(f"[flow] serving {cls.__name__} on cron {expr!r}; next slot {cron.next_after(dt.datetime.now() + offset):%Y-%m-%d %H:%M}", flush=True)
        while not max_runs or runs < max_runs:
            t = (dt.datetime.now() + offset).replace(second=0, microsecond=0)
            if t != last_slot and cron.matches(t):
                last_slot = t
                print(f"[flow] cron slot {t:%Y-%m-%d %H:%M} -> running {cls.__name__}", flush=True)
                run_id = type(self)(use_cli=False)._execute(dict(params))
                self._fire_downstream(Run(f"{cls.__name__}/{run_id}"))
                runs += 1
                continue
            time.sleep(poll_s)

    @classmethod
    def _fire_downstream(cls, finished: Run):
        """Run every deployed flow whose ``@trigger_on_finish(flow=...)`` names the flow that just finished."""
        d = os.path.join(datastore_root(), "_deployments")
        if not os.path.isdir(d):
            return
        for name in sorted(os.listdir(d)):
            with open(os.path.join(d, name)) as f:
                rec = json.load(f)
            if rec.get("decorators", {}).get("trigger_on_finish", {}).get("flow") != cls.__name__:
                continue
            spec = importlib.util.spec_from_file_location(f"_rtdc_flow_{rec['flow']}", rec["file"])
            mod = importlib.util.module_from_spec(spec)
            sys.path.insert(0, os.path.dirname(rec["file"]))
            try:
                spec.loader.exec_module(mod)
            finally:
                sys.path.pop(0)
            flow_cls = getattr(mod, rec["flow"])
            print(f"[flow] {finished.pathspec} finished -> triggering {rec['flow']}", flush=True)
            inst = flow_cls(use_cli=False)
            inst._execute({}, trigger_run=finished)


class CronSchedule:
    """Five-field cron expressions (minute hour day-of-month month day-of-week) with ``*``, lists, ranges and steps —
    what ``@schedule(cron=...)`` carries (ref train_flow.py:20). Day-of-week 0 and 7 are Sunday; when both day fields are
    restricted either may match, as in cron."""

    _RANGES = ((0, 59), (0, 23), (1, 31), (1, 12), (0, 7))

    def __init__(self, expr: str):
        parts = expr.split()
        if len(parts) != 5:
            raise ValueError(f"cron expression needs 5 fields (got {expr!r})")
        self.expr = expr
        self.fields = [self._parse(p, lo, hi) for p, (lo, hi) in zip(parts, self._RANGES)]
        self.dom_star, self.dow_star = parts[2] == "*", parts[4] == "*"
        if 7 in self.fields[4]:
            self.fields[4].add(0)

    @staticmethod
    def _parse(field: str, lo: int, hi: int) -> set:
        out = set()
        for part in field.split(","):
            step = 1
            if "/" in part:
                part, st = part.split("/", 1)
                step = int(st)
            if part in ("*", ""):
                a, b = lo, hi
            elif "-" in part:
                a, b = (int(x) for x in part.split("-", 1))
            else:
                a = int(part)
                b = hi if step > 1 else a
            if a < lo or b > hi or a > b or step < 1:
                raise ValueError(f"cron field {field!r} out of range {lo}-{hi}")
            out.update(range(a, b + 1, step))
        return out

    def matches(self, t) -> bool:
        mi, ho, dom, mo, dow = self.fields
        if t.minute not in mi or t.hour not in ho or t.month not in mo:
            return False
        d_ok, w_ok = t.day in dom, (t.isoweekday() % 7) in dow
        if self.dom_star or self.dow_star:
            return d_ok and w_ok
        return d_ok or w_ok

    def _day_matches(self, d) -> bool:
        _, _, dom, mo, dow = self.fields
        if d.month not in mo:
            return False
        d_ok, w_ok = d.day in dom, (d.isoweekday() % 7) in dow
        return (d_ok and w_ok) if (self.dom_star or self.dow_star) else (d_ok or w_ok)

    def next_after(self, t):
        """First matching minute strictly after ``t``. Walks days, not minutes (``0 0 29 2 *`` is up to eight years away:
        1896 -> 1904 skips the non-leap year 1900)."""
        import datetime as dt
        t = t.replace(second=0, microsecond=0) + dt.timedelta(minutes=1)
        minutes, hours = sorted(self.fields[0]), sorted(self.fields[1])
        day = t.replace(hour=0, minute=0)
        for i in range(366 * 8 + 2):
            if self._day_matches(day):
                for h in hours:
                    if i == 0 and h < t.hour:
                        continue
                    for m in minutes:
                        if i == 0 and h == t.hour and m < t.minute:
                            continue
                        return day.replace(hour=h, minute=m)
            day += dt.timedelta(days=1)
        raise ValueError(f"cron expression {self.expr!r} never fires")


def run_flow(flow_cls, trigger_run: Optional[Run] = None, **params) -> Run:
    """Programmatic entry: execute ``flow_cls`` with parameter overrides, return its ``Run``."""
    inst = flow_cls(use_cli=False)
    run_id = inst._execute(params, trigger_run=trigger_run)
    return Run(f"{flow_cls.__name__}/{run_id}")
