"""The FastAPI mirror of the reference's API surface (server.py:116-210): CPU tests cover the
role gates, schemas, defaults, env contract and guards (none of which touch the GPU); GPU tests
check the routes' numbers against the fixtures captured from the reference's own routes."""
import json
import os

import numpy as np
import pytest
import torch
from fastapi.testclient import TestClient

import llm_sharding_demo_b200 as L
from llm_sharding_demo_b200 import server as S

TINY = L.PRESETS["tiny-gpt2"]


def app_for(role, **kw):
    env = {"SHARD_ROLE": role, "MODEL_ID": "sshleifer/tiny-gpt2", "SPLIT_AT": "1", "GP_RANDOM_INIT": "1"}
    env.update(kw.pop("env", {}))
    return S.create_app(env=env, tokenizer=S.ByteTokenizer(), **kw)


def test_env_contract_defaults_match_reference():
    cfg = S.read_env({})
    assert cfg["MODEL_ID"] == "sshleifer/tiny-gpt2" and cfg["SHARD_ROLE"] == "coordinator"   # server.py:20-21
    assert cfg["SPLIT_AT"] == 1 and cfg["SHARD_PORT"] == 5000                               # :22,25
    assert cfg["SHARD_A_SERVICE"] == "llm-shard-a" and cfg["SHARD_B_SERVICE"] == "llm-shard-b"  # :23-24
    assert S.read_env({"SHARD_ROLE": "A"})["SHARD_ROLE"] == "a"                              # lower-cased :21


def test_role_gates_return_200_and_reference_strings(golden_dir):
    extra = json.loads(str(np.load(os.path.join(golden_dir, "tiny_gpt2.npz"))["extra"]))
    a, b, c = (TestClient(app_for(r)) for r in ("a", "b", "coordinator"))
    r = b.post("/forward", json={"input_ids": [1]})
    assert r.status_code == 200 and r.json() == extra["gate_a"]
    r = a.post("/forward_b", json={"hidden_states": [[[0.0, 0.0]]]})
    assert r.status_code == 200 and r.json() == extra["gate_b"]
    r = a.post("/generate", json={"prompt": "x"})
    assert r.status_code == 200 and r.json() == extra["gate_c"]
    r = c.post("/forward", json={"input_ids": [1]})
    assert r.json() == extra["gate_a"]


def test_schema_violations_are_422_and_guards_are_400():
    a, b, c = (TestClient(app_for(r)) for r in ("a", "b", "coordinator"))
    assert a.post("/forward", json={"input_ids": "no"}).status_code == 422
    assert a.post("/forward", json={}).status_code == 422
    assert c.post("/generate", json={"max_new_tokens": 3}).status_code == 422
    assert a.post("/forward", json={"input_ids": []}).status_code == 400
    assert a.post("/forward", json={"input_ids": [50257]}).status_code == 400
    assert a.post("/forward", json={"input_ids": [1] * 1025}).status_code == 400          # would index past wpe
    assert b.post("/forward_b", json={"hidden_states": [[[0.0] * 3]]}).status_code == 400  # wrong H
    assert b.post("/forward_b", json={"hidden_states": [[1.0], [[2.0]]]}).status_code == 400
    r = c.post("/generate", json={"prompt": "x" * 1020, "max_new_tokens": 20})
    assert r.status_code == 400 and "n_positions" in r.json()["detail"]


def test_generate_zero_tokens_returns_prompt_without_gpu(golden_dir):
    extra = json.loads(str(np.load(os.path.join(golden_dir, "tiny_gpt2.npz"))["extra"]))
    c = TestClient(app_for("coordinator"))
    r = c.post("/generate", json={"prompt": "Hi, ", "max_new_tokens": 0})
    assert r.status_code == 200 and r.json() == {"generated": extra["generate0"]}


def test_split_validation_rejects_the_shipped_k8s_mismatch():
    S.validate_split(L.PRESETS["gpt2"], 6)
    with pytest.raises(ValueError):
        S.create_app(env={"SHARD_ROLE": "a", "SPLIT_AT": "2", "MODEL_ID": "sshleifer/tiny-gpt2",
                          "GP_RANDOM_INIT": "1"})  # 2 of 2 layers
    with pytest.raises(ValueError):
        S.create_app(env={"MODEL_ID": "nope", "GP_RANDOM_INIT": "1"})


def test_server_refuses_to_start_on_a_random_network_by_accident(monkeypatch):
    """server.py:40-41 loads the pretrained model; without weights (no checkpoint, nothing cached, no network)
    the mirror must fail at start-up instead of answering 200 from a random init - unless GP_RANDOM_INIT=1."""
    monkeypatch.setenv("HF_HUB_OFFLINE", "1")
    with pytest.raises(RuntimeError, match="GP_RANDOM_INIT"):
        S.create_app(env={"SHARD_ROLE": "a", "MODEL_ID": "sshleifer/tiny-gpt2", "SPLIT_AT": "1"},
                     tokenizer=S.ByteTokenizer())
    app = S.create_app(env={"SHARD_ROLE": "a", "MODEL_ID": "sshleifer/tiny-gpt2", "SPLIT_AT": "1",
                            "GP_RANDOM_INIT": "1"}, tokenizer=S.ByteTokenizer())
    assert app.state.cfg["GP_RANDOM_INIT"] is True


def test_checkpoint_configs_that_change_the_arithmetic_are_rejected():
    base = {"n_layer": 2, "n_embd": 8, "n_head": 2}
    assert L.weights.dims_from_config(dict(base, n_inner=32)).n_embd == 8
    for bad in ({"n_inner": 16}, {"scale_attn_by_inverse_layer_idx": True}, {"reorder_and_upcast_attn": True},
                {"activation_function": "relu"}, {"scale_attn_weights": False}):
        with pytest.raises(ValueError):
            L.weights.dims_from_config(dict(base, **bad))


# ------------------------------------------------------------------------------------ GPU
def _golden(golden_dir, name):
    from oracle import gpt2_oracle as O
    z = np.load(os.path.join(golden_dir, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    dims = O.GPT2Dims(**meta["dims"])
    return z, meta, L.GPT2Dims(**meta["dims"]), O.init_weights(dims, meta["seed"], meta["recipe"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["tiny_gpt2_dense", "mini"])
def test_routes_match_reference_routes(golden_dir, name):
    z, meta, dims, w = _golden(golden_dir, name)
    env = {"SPLIT_AT": str(meta["split_at"]), "GP_DTYPE": "f32", "MODEL_ID": "x"}
    mk = lambda role, **kw: TestClient(S.create_app(env=dict(env, SHARD_ROLE=role), dims=dims, weights=w,
                                                    tokenizer=S.ByteTokenizer(), **kw))
    a, b = mk("a"), mk("b")
    prompt = z["prompt_ids"].tolist()
    r = a.post("/forward", json={"input_ids": prompt})
    assert r.status_code == 200
    hidden = r.json()["hidden_states"]
    np.testing.assert_allclose(np.asarray(hidden, np.float32), z["hidden_step0"], rtol=0, atol=3e-4)
    r = b.post("/forward_b", json={"hidden_states": hidden})
    logits = np.asarray(r.json()["logits"], np.float32)
    assert logits.shape == (1, len(prompt), dims.vocab_size)
    ref = z["logits_step0"]
    np.testing.assert_allclose(logits[0] if ref.shape[0] == len(prompt) else logits[0, -1:], ref, rtol=0, atol=3e-4)
    # B > 1 is accepted by /forward_b (server.py:120 is untyped) and rows stay independent
    r2 = b.post("/forward_b", json={"hidden_states": hidden + hidden})
    l2 = np.asarray(r2.json()["logits"], np.float32)
    assert l2.shape[0] == 2 and np.array_equal(l2[0], l2[1]) and np.array_equal(l2[0], logits[0])

    # coordinator: device loop and the reference's HTTP protocol give the same greedy text
    class Resp:
        def __init__(self, r):
            self.r = r

        def raise_for_status(self):
            self.r.raise_for_status()

        def json(self):
            return self.r.json()

    def post(url, json=None, timeout=None):
        return Resp((a if url.endswith("/forward") else b).post("/" + url.rsplit("/", 1)[1], json=json))

    n_new = 5
    text_prompt = bytes(prompt).decode("latin-1") if max(prompt) < 128 else None
    if text_prompt is not None:
        c_dev = mk("coordinator")
        c_http = TestClient(S.create_app(env=dict(env, SHARD_ROLE="coordinator", GP_COORD_MODE="http"), dims=dims,
                                         weights=w, tokenizer=S.ByteTokenizer(), http_post=post))
        g1 = c_dev.post("/generate", json={"prompt": text_prompt, "max_new_tokens": n_new}).json()["generated"]
        g2 = c_http.post("/generate", json={"prompt": text_prompt, "max_new_tokens": n_new}).json()["generated"]
        ref_ids = z["greedy_ids"][:len(prompt) + n_new].tolist()
        assert g1 == g2 == S.ByteTokenizer().decode(ref_ids)
        # default max_new_tokens is 20 (server.py:124)
        g3 = c_dev.post("/generate", json={"prompt": text_prompt}).json()["generated"]
        assert len(S.ByteTokenizer().encode(g3)) >= len(prompt)


def test_sampling_env_contract():
    assert S.read_env({})["GP_SAMPLING"] == "greedy" and S.read_env({})["GP_SAMPLE_SEED"] == 0
    assert S.read_env({"GP_SAMPLING": "Reference", "GP_SAMPLE_SEED": "7"})["GP_SAMPLING"] == "reference"
    assert (S.REF_TEMPERATURE, S.REF_TOP_K) == (0.6, 40)                                     # server.py:187-188
    with pytest.raises(ValueError):
        app_for("coordinator", env={"GP_SAMPLING": "nucleus"})


@pytest.mark.gpu
def test_generate_with_the_reference_sampler_is_seeded_and_replayable(golden_dir):
    """GP_SAMPLING=reference: /generate draws like server.py:187-205; request n of a process uses seed
    GP_SAMPLE_SEED + n, so two fresh servers replay the same answers and successive requests differ."""
    z, meta, dims, w = _golden(golden_dir, "mini")
    env = {"SPLIT_AT": str(meta["split_at"]), "GP_DTYPE": "f32", "MODEL_ID": "x", "SHARD_ROLE": "coordinator"}
    mk = lambda **e: TestClient(S.create_app(env=dict(env, **e), dims=dims, weights=w, tokenizer=S.ByteTokenizer()))
    body = {"prompt": "Hi, ", "max_new_tokens": 16}
    s1, s2 = mk(GP_SAMPLING="reference", GP_SAMPLE_SEED="5"), mk(GP_SAMPLING="reference", GP_SAMPLE_SEED="5")
    first = s1.post("/generate", json=body).json()["generated"]
    second = s1.post("/generate", json=body).json()["generated"]
    assert first == s2.post("/generate", json=body).json()["generated"]
    assert second == s2.post("/generate", json=body).json()["generated"]
    assert first != second
    greedy = mk().post("/generate", json=body).json()["generated"]
    assert greedy != first and first.startswith("Hi, ")
    assert mk(GP_SAMPLING="reference", GP_SAMPLE_SEED="6").post("/generate", json=body).json()["generated"] == second


def _write_checkpoint(tmp_path, dims, w, as_safetensors):
    """A Hugging Face style model directory built by hand: config.json + weights with the `transformer.` prefix,
    the tied lm_head and (like old GPT-2 exports) a boolean causal-mask buffer that is not a weight."""
    import struct
    d = tmp_path / ("ckpt_st" if as_safetensors else "ckpt_bin")
    d.mkdir()
    (d / "config.json").write_text(json.dumps({"n_layer": dims.n_layer, "n_embd": dims.n_embd, "n_head": dims.n_head,
                                               "vocab_size": dims.vocab_size, "n_positions": dims.n_positions,
                                               "layer_norm_epsilon": dims.layer_norm_epsilon,
                                               "activation_function": "gelu_new", "model_type": "gpt2"}))
    sd = {"transformer." + k: v for k, v in w.items()}
    sd["lm_head.weight"] = w["wte.weight"]
    if not as_safetensors:
        sd["transformer.h.0.attn.bias"] = torch.ones(1, 1, 4, 4, dtype=torch.bool)
        torch.save(sd, d / "pytorch_model.bin")
        return str(d)
    header, blobs, off = {}, [], 0
    for k, v in sd.items():
        raw = v.contiguous().numpy().astype(np.float32).tobytes()
        header[k] = {"dtype": "F32", "shape": list(v.shape), "data_offsets": [off, off + len(raw)]}
        blobs.append(raw)
        off += len(raw)
    header["transformer.h.0.attn.bias"] = {"dtype": "BOOL", "shape": [1], "data_offsets": [off, off + 1]}
    blobs.append(b"\x01")
    hj = json.dumps(header).encode()
    with open(d / "model.safetensors", "wb") as f:
        f.write(struct.pack("<Q", len(hj)) + hj + b"".join(blobs))
    return str(d)


@pytest.mark.parametrize("as_safetensors", [True, False])
def test_checkpoint_loader_reads_hf_layouts(tmp_path, as_safetensors):
    """a Hugging Face GPT-2 checkpoint (directory, safetensors or torch state_dict) yields
    the shapes and exactly the tensors the engine loader wants; nothing is downloaded (server.py:40-41 does)."""
    dims = L.PRESETS["tiny-gpt2"]
    w = L.init_weights(dims, 3, "hf")
    path = _write_checkpoint(tmp_path, dims, w, as_safetensors)
    got_dims, sd = L.load_checkpoint(path)
    assert got_dims == dims
    for name in L.stage_weight_names(dims, 0, dims.n_layer, True, True):
        assert torch.equal(sd["transformer." + name], w[name]), name
    assert all(v.dtype == torch.float32 for v in sd.values()) and "transformer.h.0.attn.bias" not in sd
    if as_safetensors:   # the hand-rolled reader agrees with the safetensors package where that is installed
        st = pytest.importorskip("safetensors.torch")
        ref = st.load_file(os.path.join(path, "model.safetensors"))
        assert all(torch.equal(ref[k], sd[k]) for k in sd)


@pytest.mark.gpu
def test_server_serves_a_checkpoint_from_disk(tmp_path, golden_dir):
    z, meta, dims, w = _golden(golden_dir, "mini")
    path = _write_checkpoint(tmp_path, dims, w, True)
    env = {"SPLIT_AT": str(meta["split_at"]), "GP_DTYPE": "f32", "SHARD_ROLE": "coordinator"}
    from_disk = TestClient(S.create_app(env=dict(env, MODEL_ID=path), tokenizer=S.ByteTokenizer()))
    from_dict = TestClient(S.create_app(env=dict(env, MODEL_ID="x"), dims=dims, weights=w, tokenizer=S.ByteTokenizer()))
    body = {"prompt": "Hi, ", "max_new_tokens": 8}
    assert from_disk.post("/generate", json=body).json() == from_dict.post("/generate", json=body).json()


@pytest.mark.gpu
def test_overlapping_generate_calls_are_batched_and_equal_their_solo_runs():
    """the reference serves one /generate at a time (server.py:154-169).  Here requests that
    overlap are decoded together, one persistent-kernel step for all of them, joining and leaving between steps -
    and every request must still get exactly the text of its solo run."""
    import threading
    from oracle import gpt2_oracle as O
    dims_o = O.PRESETS["mini"]
    dims = L.GPT2Dims(**dims_o.to_dict())
    w = O.init_weights(dims_o, 88, "dense")
    env = {"SHARD_ROLE": "coordinator", "MODEL_ID": "x", "SPLIT_AT": "2", "GP_DTYPE": "bf16"}
    batched = S.create_app(env=dict(env, GP_MAX_BATCH="4"), dims=dims, weights=w, tokenizer=S.ByteTokenizer())
    solo = TestClient(S.create_app(env=dict(env, GP_MAX_BATCH="1"), dims=dims, weights=w, tokenizer=S.ByteTokenizer()))
    client = TestClient(batched)
    reqs = [("hello there", 40), ("A", 25), ("pipeline stages", 33), ("xyz!", 1), ("the quick brown fox", 18)]
    want = [solo.post("/generate", json={"prompt": p, "max_new_tokens": n}).json()["generated"] for p, n in reqs]
    got = [None] * len(reqs)

    def call(i):
        p, n = reqs[i]
        r = client.post("/generate", json={"prompt": p, "max_new_tokens": n})
        assert r.status_code == 200, r.text
        got[i] = r.json()["generated"]
    threads = [threading.Thread(target=call, args=(i,)) for i in range(len(reqs))]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert got == want
    sched = batched.state.gp["scheduler"]
    assert sched and sched.max_joined >= 2, "the requests never shared a decode step"
    assert sorted(sched.free) == [0, 1, 2, 3] and not sched.active      # every KV slot was given back
    # a sixth request after the burst still works (slots are reusable) and 0 tokens returns the prompt
    r = client.post("/generate", json={"prompt": "again", "max_new_tokens": 3})
    assert r.json()["generated"] == solo.post("/generate", json={"prompt": "again", "max_new_tokens": 3}).json()["generated"]
    sched.close()


@pytest.mark.gpu
def test_checkpoint_on_disk_gives_the_oracle_logits(tmp_path):
    """a Hugging Face-layout GPT-2 checkpoint (config.json + model.safetensors with the
    `transformer.` prefix, Conv1D [in,out] matrices, tied lm_head) loaded by `weights.load_checkpoint` into the engine
    must give the logits the ORACLE computes from the same file read independently (server.py:40-41 is the reference
    side: from_pretrained)."""
    import struct
    from oracle import gpt2_oracle as O
    dims_o = O.PRESETS["mini"]
    w = O.init_weights(dims_o, 123, "dense")
    sd = O.hf_state_dict(w)
    # write the checkpoint by hand: safetensors container + config.json
    header, blobs, off = {}, [], 0
    for k, v in sd.items():
        b = v.contiguous().numpy().astype(np.float32).tobytes()
        header[k] = {"dtype": "F32", "shape": list(v.shape), "data_offsets": [off, off + len(b)]}
        blobs.append(b)
        off += len(b)
    hj = json.dumps(header).encode()
    d = tmp_path / "ckpt"
    d.mkdir()
    with open(d / "model.safetensors", "wb") as f:
        f.write(struct.pack("<Q", len(hj)) + hj + b"".join(blobs))
    with open(d / "config.json", "w") as f:
        json.dump({"n_layer": dims_o.n_layer, "n_embd": dims_o.n_embd, "n_head": dims_o.n_head, "vocab_size": dims_o.vocab_size,
                   "n_positions": dims_o.n_positions, "activation_function": "gelu_new", "layer_norm_epsilon": 1e-5}, f)
    ck_dims, ck_w = L.load_checkpoint(str(d))
    assert ck_dims == L.GPT2Dims(**dims_o.to_dict())
    ids = torch.randint(0, dims_o.vocab_size, (2, 11), generator=torch.Generator().manual_seed(5))
    # the oracle reads the file on its own: strip the prefix, drop the tied head
    raw = L.weights._read_safetensors(str(d / "model.safetensors"))
    ow = {k[len("transformer."):]: v for k, v in raw.items() if k.startswith("transformer.")}
    ref = O.stage_forward(ow, dims_o, ids, 0, dims_o.n_layer, True, True, None, 0)      # logits [2, 11, V]
    for dtype, tol in ((L.GP_F32, 3e-4), (L.GP_BF16, 0.06)):
        eng = L.StageEngine(ck_dims, dtype=dtype, max_batch=2)
        eng.load_weights(ck_w)
        got = eng.prefill(ids.to("cuda:0", torch.int32), use_cache=False, want_logits="all")["logits"].cpu()
        assert float((got - ref).abs().max()) <= tol, dtype
        eng.close()
