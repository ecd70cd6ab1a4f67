"""ORACLE -- test infrastructure only.

CPU restatement of the reference's torch_native attention path and KV-pool store rules.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package, and only as the checker / reported baseline; nothing under sglang_b200/
imports it.
"""
