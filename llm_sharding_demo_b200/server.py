"""FastAPI mirror of the reference's `server.py` API surface on top of the CUDA engine.

Same three routes, request bodies, defaults, env vars and role-gate behaviour:
  POST /forward    {"input_ids":[int]}            -> {"hidden_states":[[[float]*H]*T]}   (server.py:132-140)
  POST /forward_b  {"hidden_states":[B][T][H]}    -> {"logits":[B][T][V]}                (server.py:143-151)
  POST /generate   {"prompt":str,"max_new_tokens":int=20} -> {"generated":str}           (server.py:154-210)
Env: MODEL_ID, SHARD_ROLE (a|b|coordinator), SPLIT_AT, SHARD_A_SERVICE, SHARD_B_SERVICE, SHARD_PORT
(server.py:20-25).  New optional env: GP_DTYPE (bf16|f32), GP_SEED, GP_COORD_MODE (device|http),
GP_SAMPLING (greedy|reference), GP_SAMPLE_SEED, GP_WEIGHTS (checkpoint path), GP_RANDOM_INIT (0|1),
GP_MAX_BATCH (KV slots = requests decoded together).

Differences, all deliberate:
  * selection is greedy by default (BASELINE.json north_star); GP_SAMPLING=reference restores the
    reference's temperature 0.6 / top-k 40 / multinomial selection (server.py:187-205), seeded: request n
    of a process draws with seed GP_SAMPLE_SEED + n, so a restarted server replays its answers;
  * a coordinator that has a GPU runs the whole loop on the device (`gp_generate`); with
    GP_COORD_MODE=http it drives /forward and /forward_b of remote shards exactly like the
    reference, so unmodified reference shards or coordinators interoperate;
  * sequences longer than n_positions are refused with a clear error (the reference would
    raise IndexError -> 500);
  * weights, in this order: a `weights` dict passed to `create_app`; GP_WEIGHTS (or a MODEL_ID that is a local
    directory) naming a Hugging Face GPT-2 checkpoint - directory with config.json, `.safetensors` or torch
    state_dict; `AutoModelForCausalLM.from_pretrained(MODEL_ID)` exactly like server.py:40-41 (local cache
    first, then the hub).  If none of them yields weights the server REFUSES TO START - it never answers
    from a random network by accident.  GP_RANDOM_INIT=1 opts into a seeded random init of the model's
    GPT2Config shape (benchmarks, tests, boxes without checkpoints) and says so on stderr;
  * the GPT-2 BPE files come with the checkpoint; when they are missing the byte-level stand-in tokenizer is
    used and a warning is printed (token ids then differ from the reference's).
"""
from __future__ import annotations

import os
import threading
from typing import Dict, Optional

import torch
from pydantic import BaseModel

from . import weights as W
from ._lib import GP_BF16, GP_F32


# request bodies: same names, fields and default as server.py:116-124
class InputIDs(BaseModel):
    input_ids: list[int]


class HiddenStates(BaseModel):
    hidden_states: list  # nested lists (batch, seq, hidden_dim); untyped like server.py:120


class GenerateReq(BaseModel):
    prompt: str
    max_new_tokens: int = 20


class ByteTokenizer:
    """Offline stand-in when the GPT-2 BPE files are unavailable: ids are UTF-8 bytes."""

    def encode(self, text):
        return list(text.encode("utf-8"))

    def decode(self, ids, skip_special_tokens=True):
        return bytes(int(i) & 0xFF for i in ids).decode("utf-8", errors="replace")


def _warn(msg: str):
    import sys
    print(f"[llm_sharding_demo_b200] WARNING: {msg}", file=sys.stderr, flush=True)


def load_tokenizer(model_id: str):
    try:  # tokenizer stays host-side, as in server.py:40,160,209
        from transformers import AutoTokenizer
        return AutoTokenizer.from_pretrained(model_id, local_files_only=True)
    except Exception as ex:
        _warn(f"no tokenizer files for {model_id!r} ({type(ex).__name__}); using the byte-level stand-in - "
              "token ids differ from the reference's GPT-2 BPE")
        return ByteTokenizer()


def load_pretrained(model_id: str):
    """server.py:40-41: the hub model's weights, from the local cache first, then online.
    Returns (dims, state_dict) or raises."""
    from transformers import AutoModelForCausalLM
    last = None
    for kw in ({"local_files_only": True}, {}):
        if not kw and os.environ.get("HF_HUB_OFFLINE", "") == "1":
            break
        try:
            model = AutoModelForCausalLM.from_pretrained(model_id, **kw)
            dims = W.dims_from_config(model.config.to_dict())
            sd = {k: v.detach().to(torch.float32) for k, v in model.state_dict().items() if v.is_floating_point()}
            return dims, sd
        except Exception as ex:   # not cached / no network / not a GPT-2
            last = ex
    raise RuntimeError(f"cannot load pretrained weights for {model_id!r}: {last}")


def read_env(env=None) -> Dict:
    env = os.environ if env is None else env
    return {
        "MODEL_ID": env.get("MODEL_ID", "sshleifer/tiny-gpt2"),
        "SHARD_ROLE": env.get("SHARD_ROLE", "coordinator").lower(),
        "SPLIT_AT": int(env.get("SPLIT_AT", "1")),
        "SHARD_A_SERVICE": env.get("SHARD_A_SERVICE", "llm-shard-a"),
        "SHARD_B_SERVICE": env.get("SHARD_B_SERVICE", "llm-shard-b"),
        "SHARD_PORT": int(env.get("SHARD_PORT", "5000")),
        "GP_DTYPE": env.get("GP_DTYPE", "bf16"),
        "GP_SEED": int(env.get("GP_SEED", "0")),
        "GP_COORD_MODE": env.get("GP_COORD_MODE", "device"),
        "GP_SAMPLING": env.get("GP_SAMPLING", "greedy").lower(),
        "GP_SAMPLE_SEED": int(env.get("GP_SAMPLE_SEED", "0")),
        "GP_WEIGHTS": env.get("GP_WEIGHTS", ""),
        "GP_RANDOM_INIT": env.get("GP_RANDOM_INIT", "0") == "1",
        "GP_MAX_BATCH": int(env.get("GP_MAX_BATCH", "4")),
    }


REF_TEMPERATURE, REF_TOP_K = 0.6, 40   # server.py:187-188


class BatchScheduler:
    """Continuous batching for the coordinator.  The reference serves one request per loop
    (server.py:154-169: a sync handler that runs its n forward round trips to completion); here one scheduler thread
    owns the engine and advances ALL admitted requests by one token per decode step (`gp_stage_decode_slots`: one
    persistent-kernel launch for up to 8 sequences living in arbitrary KV slots).  A request joins between two steps -
    its prompt is prefilled into a free slot - and leaves when it has its max_new_tokens; the others never wait for
    it.  Greedy selection only (what the batched kernel fuses); each request's tokens are those of its solo run."""

    def __init__(self, engine, max_batch: int):
        import collections
        self.eng = engine
        self.max_batch = min(int(max_batch), 8)
        self.free = list(range(self.max_batch))
        self.active = {}                      # slot -> request dict
        self.pending = collections.deque()
        self.cv = threading.Condition()
        self.stop = False
        self.steps = 0
        self.max_joined = 0                   # most requests ever decoded in one step (observability / tests)
        self.thread = threading.Thread(target=self._run, name="gp-batch-scheduler", daemon=True)
        self.thread.start()

    def submit(self, ids, n_new):
        req = {"ids": list(ids), "n_new": int(n_new), "out": list(ids), "done": threading.Event(), "error": None}
        with self.cv:
            self.pending.append(req)
            self.cv.notify()
        req["done"].wait()
        if req["error"] is not None:
            raise req["error"]
        return req["out"]

    def close(self):
        with self.cv:
            self.stop = True
            self.cv.notify()
        self.thread.join(timeout=5)

    def _admit(self, req):
        slot = self.free.pop()
        eng = self.eng
        ids = torch.tensor([req["ids"]], dtype=torch.int32, device=eng.device)
        eng.kv_reset(slot)
        out = eng.prefill(ids, slot0=slot, use_cache=True, want_ids=True)
        tok = int(out["next_ids"][0])
        req["out"].append(tok)
        req["last"] = tok
        self.active[slot] = req

    def _retire_finished(self):
        for slot in [s for s, r in self.active.items() if len(r["out"]) - len(r["ids"]) >= r["n_new"]]:
            req = self.active.pop(slot)
            self.free.append(slot)
            req["done"].set()

    def _run(self):
        torch.cuda.set_device(self.eng.device)
        while True:
            with self.cv:
                while not self.stop and not self.pending and not self.active:
                    self.cv.wait()
                if self.stop:
                    for r in list(self.pending) + list(self.active.values()):
                        r["error"] = RuntimeError("server shutting down")
                        r["done"].set()
                    return
                joining = []
                while self.pending and len(self.free) > len(joining):
                    joining.append(self.pending.popleft())
            try:
                for req in joining:
                    self._admit(req)
                self._retire_finished()          # max_new_tokens == 1: the prefill already produced it
                if self.active:
                    slots = sorted(self.active)
                    toks = torch.tensor([self.active[s]["last"] for s in slots], dtype=torch.int32, device=self.eng.device)
                    nxt = self.eng.decode_slots(toks, slots)["next_ids"].cpu().tolist()
                    self.steps += 1
                    self.max_joined = max(self.max_joined, len(slots))
                    for s, t in zip(slots, nxt):
                        self.active[s]["out"].append(int(t))
                        self.active[s]["last"] = int(t)
                    self._retire_finished()
            except Exception as ex:   # a failed step fails the requests it carried, not the server
                for r in joining + list(self.active.values()):
                    if not r["done"].is_set():
                        r["error"] = ex
                        r["done"].set()
                for s in list(self.active):
                    self.free.append(s)
                self.active.clear()


def validate_split(dims: W.GPT2Dims, split_at: int):
    """The reference ships SPLIT_AT=2 on shard A and 1 on shard B (k8s/*-deployment.yaml:22-23),
    which silently runs block 1 twice; here the split must leave both shards non-empty."""
    if not 0 < split_at < dims.n_layer:
        raise ValueError(f"SPLIT_AT={split_at} must satisfy 0 < SPLIT_AT < n_layer={dims.n_layer}")


def create_app(env=None, dims: Optional[W.GPT2Dims] = None, weights=None, tokenizer=None, http_post=None):
    """Build the FastAPI app for one role.  `dims`/`weights`/`tokenizer`/`http_post` are
    injection points for tests; production use reads everything from the environment."""
    from fastapi import FastAPI, HTTPException

    cfg = read_env(env)
    role = cfg["SHARD_ROLE"]
    ckpt = cfg["GP_WEIGHTS"] or (cfg["MODEL_ID"] if os.path.isdir(cfg["MODEL_ID"]) else "")
    if ckpt and weights is None:
        ck_dims, weights = W.load_checkpoint(ckpt)
        dims = dims or ck_dims
    if weights is None and not cfg["GP_RANDOM_INIT"]:
        try:
            hub_dims, weights = load_pretrained(cfg["MODEL_ID"])
            dims = dims or hub_dims
        except Exception as ex:
            raise RuntimeError(
                f"{ex}.  Point GP_WEIGHTS (or MODEL_ID) at a local GPT-2 checkpoint, or set GP_RANDOM_INIT=1 to "
                "serve a seeded random-init network of that shape on purpose.") from ex
    dims = dims or W.PRESETS.get(cfg["MODEL_ID"])
    if dims is None:
        raise ValueError(f"unknown MODEL_ID {cfg['MODEL_ID']!r}; known: {sorted(W.PRESETS)}")
    if weights is None:
        _warn(f"GP_RANDOM_INIT=1: serving a RANDOM-INIT {cfg['MODEL_ID']} (seed {cfg['GP_SEED']}); "
              "outputs are not the pretrained model's")
    validate_split(dims, cfg["SPLIT_AT"])
    tokenizer = tokenizer or load_tokenizer(cfg["MODEL_ID"])
    dtype = GP_BF16 if cfg["GP_DTYPE"] == "bf16" else GP_F32
    print("Starting server role:", role, "model:", cfg["MODEL_ID"])

    if cfg["GP_SAMPLING"] not in ("greedy", "reference"):
        raise ValueError(f"GP_SAMPLING={cfg['GP_SAMPLING']!r} must be 'greedy' or 'reference'")
    sampled = cfg["GP_SAMPLING"] == "reference"
    lock = threading.Lock()   # an engine handle is not re-entrant; handlers run on a thread pool
    state = {"engine": None, "requests": 0, "scheduler": None}
    max_batch = max(1, min(8, cfg["GP_MAX_BATCH"])) if role == "coordinator" else 1

    def engine():
        if state["engine"] is None:
            from .engine import StageEngine
            split = cfg["SPLIT_AT"]
            if role == "a":
                rng = (0, split, True, False)
            elif role == "b":
                rng = (split, dims.n_layer, False, True)
            else:
                rng = (0, dims.n_layer, True, True)
            eng = StageEngine(dims, rng[0], rng[1], has_embed=rng[2], has_head=rng[3], dtype=dtype, max_batch=max_batch)
            names = eng.weight_names()
            w = weights if weights is not None else W.init_weights(dims, cfg["GP_SEED"], "hf", only=names)
            eng.load_weights(w)
            state["engine"] = eng
        return state["engine"]

    def scheduler():
        """The continuous-batching scheduler, when this coordinator can use it: greedy selection on the bf16 persistent
        kernel (head_dim 64, hidden size a multiple of 64) with more than one KV slot; else None (one request at a time)."""
        if state["scheduler"] is None:
            ok = (max_batch > 1 and not sampled and dtype == GP_BF16 and dims.head_dim == 64 and dims.n_embd % 64 == 0)
            state["scheduler"] = BatchScheduler(engine(), max_batch) if ok else False
        return state["scheduler"] or None

    app = FastAPI()
    app.state.cfg = cfg
    app.state.gp = state

    def bad(msg):
        return HTTPException(status_code=400, detail=msg)

    def check_ids(ids):
        if len(ids) == 0:
            raise bad("input_ids is empty")
        if len(ids) > dims.n_positions:
            raise bad(f"sequence of {len(ids)} tokens exceeds n_positions={dims.n_positions}")
        if min(ids) < 0 or max(ids) >= dims.vocab_size:
            raise bad(f"token id outside [0,{dims.vocab_size})")

    @app.post("/forward")
    def forward_a(req: InputIDs):
        if role != "a":
            return {"error": "This instance is not shard A."}
        check_ids(req.input_ids)
        with lock:
            eng = engine()
            ids = torch.tensor([req.input_ids], dtype=torch.int32, device=eng.device)
            hidden = eng.prefill(ids, use_cache=False)["hidden"]
            return {"hidden_states": hidden.cpu().tolist()}

    @app.post("/forward_b")
    def forward_b(req: HiddenStates):
        if role != "b":
            return {"error": "This instance is not shard B."}
        try:
            hidden = torch.tensor(req.hidden_states, dtype=torch.float32)
        except (ValueError, TypeError) as ex:
            raise bad(f"hidden_states is not a rectangular nested list of numbers: {ex}")
        if hidden.dim() != 3 or hidden.shape[2] != dims.n_embd or hidden.shape[1] == 0:
            raise bad(f"hidden_states must be [batch, seq, {dims.n_embd}]")
        if hidden.shape[1] > dims.n_positions:
            raise bad(f"sequence of {hidden.shape[1]} tokens exceeds n_positions={dims.n_positions}")
        with lock:
            eng = engine()
            outs = []
            for b in range(hidden.shape[0]):   # one sequence per call keeps the workspace at T*V
                x = hidden[b:b + 1].to(eng.device)
                outs.append(eng.prefill(x, use_cache=False, want_logits="all")["logits"].cpu())
            return {"logits": torch.cat(outs, 0).tolist()}

    @app.post("/generate")
    def generate(req: GenerateReq):
        if role != "coordinator":
            return {"error": "This instance is not coordinator."}
        input_ids = tokenizer.encode(req.prompt)
        n_new = max(0, int(req.max_new_tokens))
        if n_new == 0:
            return {"generated": tokenizer.decode(input_ids, skip_special_tokens=True)}
        check_ids(input_ids)
        if len(input_ids) + n_new - 1 > dims.n_positions:
            raise bad(f"prompt of {len(input_ids)} tokens + {n_new} new tokens exceeds "
                      f"n_positions={dims.n_positions}")
        if cfg["GP_COORD_MODE"] == "http":
            # the reference's wire protocol (server.py:169-206), greedy selection
            import requests
            post = http_post or requests.post
            ids = list(input_ids)
            with lock:
                state["requests"] += 1
                gen = torch.Generator().manual_seed(cfg["GP_SAMPLE_SEED"] + state["requests"] - 1)
            for _ in range(n_new):
                url_a = f"http://{cfg['SHARD_A_SERVICE']}:{cfg['SHARD_PORT']}/forward"
                resp = post(url_a, json={"input_ids": ids}, timeout=30)
                resp.raise_for_status()
                hidden = resp.json()["hidden_states"]
                url_b = f"http://{cfg['SHARD_B_SERVICE']}:{cfg['SHARD_PORT']}/forward_b"
                resp2 = post(url_b, json={"hidden_states": hidden}, timeout=30)
                resp2.raise_for_status()
                last = torch.tensor(resp2.json()["logits"][0][-1], dtype=torch.float32)
                if sampled:   # server.py:190-205 verbatim in meaning, with a seeded generator
                    vals, idx = torch.topk(last / REF_TEMPERATURE, min(REF_TOP_K, last.numel()))
                    probs = torch.softmax(vals, dim=-1)
                    ids.append(int(idx[torch.multinomial(probs, 1, generator=gen)]))
                else:
                    ids.append(int(torch.argmax(last)))
            return {"generated": tokenizer.decode(ids, skip_special_tokens=True)}
        with lock:
            sched = scheduler()
        if sched is not None:
            try:
                out_ids = sched.submit(input_ids, n_new)
            except Exception as ex:
                raise HTTPException(status_code=500, detail=str(ex))
            return {"generated": tokenizer.decode(out_ids, skip_special_tokens=True)}
        with lock:
            eng = engine()
            state["requests"] += 1
            if sampled:
                eng.set_sampler(REF_TEMPERATURE, min(REF_TOP_K, dims.vocab_size), cfg["GP_SAMPLE_SEED"] + state["requests"] - 1)
            out = eng.generate(torch.tensor([input_ids], dtype=torch.int32), n_new)
        return {"generated": tokenizer.decode(out[0].tolist(), skip_special_tokens=True)}

    return app


def __getattr__(name):
    # `uvicorn llm_sharding_demo_b200.server:app` (Dockerfile:19 of the reference uses server:app)
    if name == "app":
        return create_app()
    raise AttributeError(name)
