"""CPU oracle of the reference's pipeline-sharded GPT-2 path.  Test infrastructure only:
nothing under `llm_sharding_demo_b200/` may import this package."""
