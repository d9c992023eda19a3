"""Operator layer of the Polya-Gamma EFGP classifier on the B200 hot path."""
