"""Seeded synthetic workloads (reference FASTA, pore model, reads, FAST5, chains) for tests/ and bench.py.
Test and benchmark infrastructure: not part of the product package rawalign_b200."""
