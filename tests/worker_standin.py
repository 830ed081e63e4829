"""TEST INFRASTRUCTURE -- restatement of the reference's Python worker for machines without /root/reference.

Follows src/native/python/python_algorithm_reply.py (the Rust server's child process): same command line
(:135-153), same loader (`sys.path.append(dir)`, digit strings -> int, `config_path` added, import
`{name}.{name}`, `cls(**hyperparams)`, `[algorithm_pyscript_true]` marker on stdout; :23-52), same JSON-lines loop
(:155-177) with every algorithm call under one asyncio.Lock and run through asyncio.to_thread (:66-112), same replies.

tests/test_worker_pipe.py checks, where the reference tree exists, that this file and the UNMODIFIED reference worker
answer identically line for line; the GPU box (no reference tree) then drives REINFORCE_B200 through this one.
"""
import argparse
import asyncio
import importlib
import json
import logging
import sys

logging.basicConfig(stream=sys.stdout, level=logging.INFO)

from relayrl_framework import RelayRLTrajectory  # noqa: E402  (tests/standin on PYTHONPATH)


def load_algorithm(name, hyperparams, algorithm_dir, config_path):
    sys.path.append(algorithm_dir)
    hp = json.loads(hyperparams) if hyperparams else None
    if hp is not None:
        for k, v in hp.items():
            if isinstance(v, str) and v.isdigit():
                hp[k] = float(v) if "." in v else int(v)
    hp["config_path"] = config_path
    cls = getattr(importlib.import_module("%s.%s" % (name, name)), name)
    algo = cls(**hp) if cls else None
    logging.critical("[algorithm_pyscript_true]" if algo is not None else "[algorithm_pyscript_false]")
    return algo


async def main():
    ap = argparse.ArgumentParser()
    for a in ("--algorithm_name", "--env_dir", "--resolved_algorithm_dir", "--config_path", "--hyperparams"):
        ap.add_argument(a, type=str, default=None)
    args = ap.parse_args()
    algo = load_algorithm(args.algorithm_name, args.hyperparams, args.resolved_algorithm_dir, args.config_path)
    lock = asyncio.Lock()

    async def handle(cmd):
        name = cmd.get("command")
        if name == "save_model":
            async with lock:
                if not algo:
                    return {"status": "error", "message": "Algorithm not initialized"}
                try:
                    await asyncio.to_thread(algo.save)
                    return {"status": "success"}
                except Exception as e:
                    return {"status": "error", "message": str(e)}
        if name == "receive_trajectory":
            traj = await asyncio.to_thread(RelayRLTrajectory.traj_from_json, cmd.get("trajectory"))
            async with lock:
                if not algo:
                    return {"status": "error", "message": "Algorithm not initialized"}
                try:
                    updated = await asyncio.to_thread(algo.receive_trajectory, traj)
                    return {"status": "success"} if updated else {"status": "not_updated"}
                except Exception as e:
                    return {"status": "error", "message": str(e)}
        return {"status": "error", "message": "Unknown command: %s" % name}

    loop = asyncio.get_running_loop()
    reader = asyncio.StreamReader(limit=10240 * 10240)
    await loop.connect_read_pipe(lambda: asyncio.StreamReaderProtocol(reader), sys.stdin)
    while True:
        line = await reader.readline()
        if not line:
            break
        try:
            resp = await handle(json.loads(line.decode().strip()))
        except json.JSONDecodeError:
            resp = {"status": "error", "message": "Invalid JSON"}
        except Exception as e:
            resp = {"status": "error", "message": str(e)}
        print(json.dumps(resp))
        sys.stdout.flush()
    print("[PythonWorker] Shutting down worker...")


if __name__ == "__main__":
    asyncio.run(main())
