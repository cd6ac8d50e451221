"""Reference-facing entry points (same module and function names as the reference's libs/)."""
