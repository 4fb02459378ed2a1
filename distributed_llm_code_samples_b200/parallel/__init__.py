"""Parallel strategies (single / DDP / FSDP / MP), the process launcher and the communication layers."""
