"""Input generators of the benchmark and the tests (no product code, no oracle arithmetic)."""
