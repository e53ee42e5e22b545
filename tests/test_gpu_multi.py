"""Runs tests/multi_gpu_check.py under torchrun when the box has >= 2 GPUs (skipped otherwise)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_pipeline_across_gpus(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29620 + world), os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MULTI_GPU_CHECK_OK" in r.stdout


def test_two_stages_in_one_process():
    """A host that owns several GPUs in ONE process connects its stages with gp_comm_connect_local (CUDA IPC handles
    cannot be opened by the exporting process): two stages on two GPUs, driven from two host threads, must produce the
    tokens of the single-stage engine."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import threading
    import llm_sharding_demo_b200 as L
    from oracle import gpt2_oracle as O
    dims = O.PRESETS["mini"]
    w = O.init_weights(dims, 71, "dense")
    pd = L.GPT2Dims(**dims.to_dict())
    s0 = L.StageEngine(pd, 0, 2, has_embed=True, has_head=False, dtype=L.GP_BF16, max_batch=2, device=0)
    s1 = L.StageEngine(pd, 2, 4, has_embed=False, has_head=True, dtype=L.GP_BF16, max_batch=2, device=1)
    single = L.StageEngine(pd, dtype=L.GP_BF16, max_batch=2, device=0)
    for e in (s0, s1, single):
        e.load_weights(w)
    s0.comm_export(); s1.comm_export()
    s0.comm_connect_local(s1); s1.comm_connect_local(s0)
    prompts = torch.randint(0, dims.vocab_size, (2, 6), generator=torch.Generator().manual_seed(3), dtype=torch.int32)
    n_new = 8
    want = single.generate(prompts, n_new)
    # prefill through both stages (host-side hop), then the device-resident decode on two threads
    with torch.cuda.device(0):
        h = s0.prefill(prompts.to("cuda:0"), use_cache=True)["hidden"]
    with torch.cuda.device(1):
        first = s1.prefill(h.to("cuda:1"), use_cache=True, want_ids=True)["next_ids"]
    first0 = first.to("cuda:0")
    res = {}

    def run(eng, dev, ids):
        with torch.cuda.device(dev):
            res[dev] = eng.pipeline_decode(ids, n_new - 1, 2, 1, comm_epoch0=1, timeout_ms=5000)
    t0 = threading.Thread(target=run, args=(s0, 0, first0))
    t1 = threading.Thread(target=run, args=(s1, 1, None))
    t0.start(); t1.start(); t0.join(); t1.join()
    tokens = res[0][0].cpu()
    got = torch.cat([prompts, first.cpu().view(2, 1), tokens], dim=1)
    assert torch.equal(got, want)


@pytest.mark.parametrize("sampling", ["greedy", "reference"])
def test_pipeline_pod_serves_generate(sampling):
    """The N-GPU pod of k8s/pipeline-pod.yaml (`torchrun -m llm_sharding_demo_b200.pod`): /healthz turns ready, and
    /generate returns the text a single-GPU engine produces for the same prompt and weights - greedy, and with the
    reference's sampler (GP_SAMPLING=reference: drawn inside the last stage's kernel, seed GP_SAMPLE_SEED + request
    number, KV slot 0 - the draw a single-stage engine makes with the same seed)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import signal
    import time
    import requests
    import llm_sharding_demo_b200 as L
    from llm_sharding_demo_b200.server import ByteTokenizer
    port = 5057
    port += sampling == "reference"
    env = dict(os.environ, MODEL_ID="gpt2", GP_RANDOM_INIT="1", GP_SEED="3", SHARD_PORT=str(port), GP_MAX_CTX="96",
               GP_SAMPLING=sampling, GP_SAMPLE_SEED="41",
               PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(29688 + (sampling == "reference")), "-m", "llm_sharding_demo_b200.pod"]
    proc = subprocess.Popen(cmd, env=env, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, start_new_session=True)
    try:
        ready = False
        for _ in range(240):
            if proc.poll() is not None:
                break
            try:
                if requests.get(f"http://127.0.0.1:{port}/healthz", timeout=1).status_code == 200:
                    ready = True
                    break
            except Exception:
                time.sleep(0.5)
        assert ready, (proc.stdout.read().decode()[-2000:] if proc.poll() is not None else "pod never became ready")
        r = requests.post(f"http://127.0.0.1:{port}/generate", json={"prompt": "Hi, ", "max_new_tokens": 12}, timeout=60)
        assert r.status_code == 200, r.text
        dims = L.PRESETS["gpt2"]
        eng = L.StageEngine(dims, dtype=L.GP_BF16, max_batch=1, max_ctx=96)
        eng.load_weights(L.init_weights(dims, 3, "hf"))
        if sampling == "reference":
            eng.set_sampler(0.6, 40, seed=41)          # first request of the pod: GP_SAMPLE_SEED + 0
        tok = ByteTokenizer()
        want = eng.generate(torch.tensor([tok.encode("Hi, ")], dtype=torch.int32), 12)
        assert r.json()["generated"] == tok.decode(want[0].tolist())
        # the reference's default and its zero-token case
        assert requests.post(f"http://127.0.0.1:{port}/generate", json={"prompt": "abc", "max_new_tokens": 0}, timeout=30).json() == {"generated": "abc"}
    finally:
        os.killpg(proc.pid, signal.SIGTERM)
        try:
            proc.wait(timeout=20)
        except Exception:
            os.killpg(proc.pid, signal.SIGKILL)
