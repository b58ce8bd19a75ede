"""Data-parallel runtime: process group facade, symmetric-memory communicator, reducer, DDP."""
