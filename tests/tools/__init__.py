"""Checker-side GPU tools (they import oracle/, which only tests/, smoke() and the bench baselines may): parity reports and stage-by-stage debuggers."""
