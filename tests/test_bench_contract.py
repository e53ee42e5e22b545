"""bench.py's host-side contract (no GPU): both arms describe the workload with the same `config` object, the
algorithmic byte count follows SURVEY.md 8(d), and the synthetic prompts are reproducible."""
import bench


def test_both_arms_share_one_config_object():
    for name in bench.WORKLOADS:
        for n in (1, 2, 4, 8):
            c = bench.workload_config(name, n)
            assert c["workload"] == name and c["streams"] == n and c["stages"] == n
            assert set(c) == {"workload", "model", "batch_per_stream", "streams", "prompt", "new_tokens", "stages", "l2"}
    assert bench.DEFAULT_WORKLOAD == "gpt2-124m-b1-p32-n128"      # BASELINE.json configs[1]


def test_algorithmic_bytes_per_step_matches_the_survey_example():
    import llm_sharding_demo_b200.weights as W
    dims = W.PRESETS["gpt2"]
    # 124M, B=1, T=160, bf16: 247.3 MB of weights + 5.9 MB of KV = 253 MB per token
    b = bench.algorithmic_bytes_per_step(dims, 1, 160, 1, 2)
    assert abs(b / 1e6 - 253.2) < 1.0
    # stage hops add 4*B*H bytes per boundary
    assert bench.algorithmic_bytes_per_step(dims, 1, 160, 8, 2) - b == 7 * 4 * 768


def test_synthetic_prompts_are_seeded():
    a = bench.synthetic_prompts(2, 32, 50257)
    assert a.shape == (2, 32) and int(a.min()) >= 0 and int(a.max()) < 50257
    assert (a == bench.synthetic_prompts(2, 32, 50257)).all()
