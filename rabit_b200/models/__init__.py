"""Workloads built on the collectives: gradient-histogram aggregation and torch checkpoint helpers."""
