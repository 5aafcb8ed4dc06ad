"""Entry point of one worker process started by ``TorchTrainer.fit()`` (driver mode).

Does what a Ray Train worker actor does before calling the user's function (implicit at
``/root/reference/my_ray_module.py:244-250``): read the rendezvous environment, join the process
group (``nccl`` when ``use_gpu`` else ``gloo``), install the session, run
``train_loop_per_worker(config)``, report completion or the traceback to the driver.
"""
from __future__ import annotations

import os
import sys
import traceback


def main(payload_path: str) -> int:
    import cloudpickle  # noqa: F401  (registers reducers used by pickle.load below)
    import pickle

    with open(payload_path, "rb") as f:
        payload = pickle.load(f)      # the driver put its import roots on our PYTHONPATH
    for p in payload.get("sys_path", []):
        if p and p not in sys.path:
            sys.path.append(p)

    from ..train.link import SocketLink
    from ..train.session import TrainContext, _Session, _set_session
    from ..train.trainer import _call_train_fn

    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    link = SocketLink(payload["socket"], payload["authkey"], rank, heartbeat_s=payload.get("heartbeat_s", 0.0))
    try:
        import torch
        import torch.distributed as dist

        use_gpu = payload["use_gpu"]
        if use_gpu:
            torch.cuda.set_device(local_rank)
        if world > 1:
            kwargs = {"device_id": torch.device("cuda", local_rank)} if use_gpu else {}
            dist.init_process_group(backend=payload["backend"], rank=rank, world_size=world, **kwargs)
        storage = payload["storage"]
        ctx = TrainContext(rank, world, local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", world)),
                           int(os.environ.get("NODE_RANK", 0)), payload["experiment_name"], payload["trial_name"],
                           storage.trial_path, storage, use_gpu, payload.get("metadata"))
        session = _Session(ctx, link, payload.get("start_checkpoint"), payload.get("start_index", 0))
        _set_session(session)
        prof_path = os.environ.get("RTDC_WORKER_CPROFILE")           # host-side profile of the training function, per rank
        if prof_path:
            import cProfile
            import pstats
            prof = cProfile.Profile()
            prof.enable()
        _call_train_fn(payload["fn"], payload["config"])
        if prof_path:
            prof.disable()
            with open(f"{prof_path}.{rank}", "w") as f:
                pstats.Stats(prof, stream=f).sort_stats("cumulative").print_stats(70)
        session.drain()
        if use_gpu:
            torch.cuda.synchronize()
        from ..data.loader import dump_timing
        dump_timing(rank)
        from ..ckpt import async_writer
        async_writer.wait_all()
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        link.done()
        return 0
    except SystemExit as e:
        return int(e.code or 0)
    except BaseException:
        tb = traceback.format_exc()
        sys.stderr.write(tb)
        link.error(tb)
        return 1
    finally:
        _set_session(None)
        link.close()


if __name__ == "__main__":
    code = main(sys.argv[1])
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(code)
