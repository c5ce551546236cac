"""Rendezvous tracker and local launcher (the part of dmlc-core the reference depends on)."""
