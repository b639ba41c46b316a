"""Multi-GPU PPO harness for the G1 velocity task on the engine."""
