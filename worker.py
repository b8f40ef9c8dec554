#!/usr/bin/env python
"""Client entrypoint - same invocation as the reference: `python worker.py [server_ip own_ip]`, then
`request live`, `request <path-to-video>`, `end`, or anything else to quit.
Optional `--flag value` pairs override Config defaults (e.g. --my-port 0 --device cuda:3)."""
from distributedvolunteercomputing_b200.client import client, main  # noqa: F401

if __name__ == "__main__":
    main()
