#!/usr/bin/env python
"""Server entrypoint - same invocation as the reference: `python server.py [ip]`, then `quit`.
Optional `--flag value` pairs override Config defaults (see distributedvolunteercomputing_b200/config.py)."""
from distributedvolunteercomputing_b200.coordinator import coordinator, main  # noqa: F401

if __name__ == "__main__":
    main()
