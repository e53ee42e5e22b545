"""Pins the CPU oracle (`oracle/gpt2_oracle.py`) against outputs of the reference's own
`server.py`, captured by `oracle/make_golden.py` into `tests/golden/` (CPU only)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import gpt2_oracle as O

SMALL = ["tiny_gpt2", "tiny_gpt2_dense", "mini", "mini32"]


def load(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    dims = O.GPT2Dims(**meta["dims"])
    w = O.init_weights(dims, meta["seed"], meta["recipe"])
    return z, meta, dims, w


@pytest.mark.parametrize("name", SMALL)
def test_shard_a_hidden_matches_reference(golden_dir, name):
    z, meta, dims, w = load(golden_dir, name)
    ids = torch.tensor(z["prompt_ids"]).unsqueeze(0)
    h = O.shard_a_forward(w, dims, ids, meta["split_at"])
    np.testing.assert_allclose(h.numpy(), z["hidden_step0"], rtol=0, atol=2e-5)


@pytest.mark.parametrize("name", SMALL)
def test_shard_b_logits_match_reference(golden_dir, name):
    z, meta, dims, w = load(golden_dir, name)
    hidden = torch.tensor(z["hidden_step0"])
    logits = O.shard_b_forward(w, dims, hidden, meta["split_at"])[0]
    ref = z["logits_step0"]
    got = logits.numpy() if ref.shape[0] == logits.shape[0] else logits[-1:].numpy()
    np.testing.assert_allclose(got, ref, rtol=0, atol=2e-5)


@pytest.mark.parametrize("name", SMALL)
def test_greedy_ids_match_reference_recompute_and_cached(golden_dir, name):
    z, meta, dims, w = load(golden_dir, name)
    prompt = torch.tensor(z["prompt_ids"]).unsqueeze(0)
    ref_ids = torch.tensor(z["greedy_ids"]).unsqueeze(0)
    n_new = ref_ids.shape[1] - prompt.shape[1]
    margins = z["margins"]
    ids, logs = O.generate_recompute(w, dims, prompt, n_new, meta["split_at"], return_logits=True)
    # bit-exact ids wherever the reference's top-2 margin exceeds 2x the fp32 tolerance
    tol = 2e-5
    for t in range(n_new):
        if margins[t] > 2 * tol:
            assert int(ids[0, prompt.shape[1] + t]) == int(ref_ids[0, prompt.shape[1] + t]), t
        else:
            break  # after a sub-tolerance margin the rollouts may legitimately diverge
    # cached decode, teacher-forced on the reference ids, 1 and 2 stages
    for S in (1, 2):
        if S > dims.n_layer:
            continue
        ids_c, logs_c = O.generate_cached(w, dims, prompt, n_new, n_stages=S, teacher=ref_ids,
                                          return_logits=True)
        for t in range(n_new):
            if margins[t] > 2 * tol:
                assert int(ids_c[0, prompt.shape[1] + t]) == int(ref_ids[0, prompt.shape[1] + t]), (S, t)
        ll = z["last_logits"]
        if ll.shape[0] == n_new:
            np.testing.assert_allclose(logs_c[0].numpy(), ll, rtol=0, atol=3e-5)
        else:
            np.testing.assert_allclose(logs_c[0, [0, n_new - 1]].numpy(), ll, rtol=0, atol=3e-5)


def test_gpt2_124m_against_reference(golden_dir):
    z, meta, dims, w = load(golden_dir, "gpt2_124m")
    prompt = torch.tensor(z["prompt_ids"]).unsqueeze(0)
    ref_ids = torch.tensor(z["greedy_ids"]).unsqueeze(0)
    n_new = 4
    h = O.shard_a_forward(w, dims, prompt, meta["split_at"])
    np.testing.assert_allclose(h[0, -1].numpy(), z["hidden_step0_last"], rtol=0, atol=5e-5)
    ids_c, logs = O.generate_cached(w, dims, prompt, n_new, n_stages=2, teacher=ref_ids, return_logits=True)
    np.testing.assert_allclose(logs[0, 0].numpy(), z["last_logits"][0], rtol=0, atol=5e-5)
    for t in range(n_new):
        assert int(ids_c[0, 32 + t]) == int(ref_ids[0, 32 + t])
        got = logs[0, t][torch.tensor(z["top8_idx"][t])].numpy()
        np.testing.assert_allclose(got, z["top8_val"][t], rtol=0, atol=5e-5)


def test_reference_api_fixture_semantics(golden_dir):
    """Route behaviour captured from the reference (role gates `server.py:134-157`, default
    max_new_tokens=20 `server.py:124`, max_new_tokens=0 returns the decoded prompt)."""
    z = np.load(os.path.join(golden_dir, "tiny_gpt2.npz"))
    extra = json.loads(str(z["extra"]))
    assert extra["gate_a"] == {"error": "This instance is not shard A."}
    assert extra["gate_b"] == {"error": "This instance is not shard B."}
    assert extra["gate_c"] == {"error": "This instance is not coordinator."}
    assert extra["generate0"] == "Hi, "
    assert extra["generate_default_len"] >= 4  # 4 prompt bytes + up to 20 sampled tokens


def test_split_layers_and_stage_names():
    assert O.split_layers(12, 2) == [(0, 6), (6, 12)]
    assert O.split_layers(48, 8)[-1] == (42, 48)
    assert O.split_layers(2, 2) == [(0, 1), (1, 2)]
    dims = O.PRESETS["mini"]
    n0 = O.stage_weight_names(dims, 0, 2, True, False)
    n1 = O.stage_weight_names(dims, 2, 4, False, True)
    assert "wpe.weight" in n0 and "wpe.weight" not in n1
    assert "wte.weight" in n0 and "wte.weight" in n1   # tied head duplicated on the last stage
    assert "ln_f.weight" in n1 and "ln_f.weight" not in n0
    w_all = O.init_weights(dims, 5, "dense")
    w_part = O.init_weights(dims, 5, "dense", only=n1)
    for k in n1:
        assert torch.equal(w_all[k], w_part[k])


def test_oracle_matches_hf_model_directly():
    """Second pin: the installed transformers GPT2LMHeadModel on the same weights."""
    from transformers import GPT2Config, GPT2LMHeadModel
    dims = O.PRESETS["mini"]
    w = O.init_weights(dims, 21, "dense")
    cfg = GPT2Config(vocab_size=dims.vocab_size, n_positions=dims.n_positions, n_embd=dims.n_embd,
                     n_layer=dims.n_layer, n_head=dims.n_head)
    m = GPT2LMHeadModel(cfg).eval()
    m.load_state_dict(O.hf_state_dict(w), strict=False)
    ids = torch.tensor([[3, 9, 27, 81, 243], [1, 1, 2, 3, 5]])
    with torch.no_grad():
        ref = m(ids).logits
    got = O.stage_forward(w, dims, ids, 0, dims.n_layer, True, True)
    assert float((ref - got).abs().max()) < 3e-5


# ----------------------------------------------------------------------- sampled selection
@pytest.mark.parametrize("name", ["sampled_mini", "sampled_tiny_gpt2"])
def test_reference_sampler_restatement_matches_reference_generate(golden_dir, name):
    """The unmodified reference's /generate, run under torch.manual_seed (oracle/make_golden.py), against the
    oracle's line-for-line restatement of server.py:169-206 under the same seed: identical token ids."""
    g = np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)
    meta = json.loads(str(g["meta"]))
    dims = O.GPT2Dims(**meta["dims"])
    w = O.init_weights(dims, meta["seed"], meta["recipe"])
    prompt = list(meta["prompt"].encode("utf-8")) if isinstance(meta["prompt"], str) else meta["prompt"]
    for ts, want in zip(g["torch_seeds"].tolist(), g["ids"].tolist()):
        got = O.generate_reference_sampled(w, dims, prompt, int(g["n_new"]), meta["split_at"], ts)
        assert got == want


def test_philox_known_answers():
    """Random123 known-answer vectors for philox4x32, 10 rounds."""
    kat = [((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
           ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
           ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
            (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1))]
    for ctr, key, want in kat:
        assert tuple(O.philox4x32_10(ctr, key)) == want


def test_seeded_sampler_uses_the_reference_candidates_and_distribution():
    gen = torch.Generator().manual_seed(3)
    logits = torch.randn(50257, generator=gen) * 2.0
    idx, probs = O.topk_probs(logits, 0.6, 40)
    vals, ref_idx = torch.topk(logits / 0.6, 40)          # server.py:196
    assert torch.equal(idx, ref_idx)
    assert torch.allclose(probs, torch.softmax(vals, -1), atol=0, rtol=0)
    # ties resolve to the lower token id, also across the k-th place
    flat = torch.zeros(100)
    flat[[7, 3, 50]] = 1.0
    assert O.topk_probs(flat, 0.6, 5)[0].tolist() == [3, 7, 50, 0, 1]
    # the seeded draw follows those probabilities (chi-square, 39 degrees of freedom, 99.9 % point = 72.1)
    n = 3000
    counts = torch.zeros(40)
    where = {int(t): j for j, t in enumerate(idx.tolist())}
    for pos in range(n):
        tok, _ = O.sample_select(logits, 0.6, 40, seed=77, row=0, position=pos)
        counts[where[tok]] += 1
    expect = probs * n
    big = expect >= 5          # pool the sparse tail as usual for the test
    chi2 = float((((counts - expect) ** 2) / expect)[big].sum())
    chi2 += float((counts[~big].sum() - expect[~big].sum()) ** 2 / max(float(expect[~big].sum()), 1e-9))
    assert chi2 < 72.1, chi2
    # and it is a pure function of (seed, row, position)
    a = O.sample_select(logits, 0.6, 40, 77, 1, 5)[0]
    assert a == O.sample_select(logits, 0.6, 40, 77, 1, 5)[0]
    draws = {O.sample_select(logits, 0.6, 40, 77, r, 5)[0] for r in range(16)}
    assert len(draws) > 1
