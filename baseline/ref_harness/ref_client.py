"""Wrapper that runs the UNMODIFIED reference `worker.py` (imported from baseline/_ref) with a unique
port and chunk size.  Both are class attributes in the reference (worker.py:16,25), so no reference
file is edited.  Replays the reference REPL (worker.py:346-356) from stdin.

    python ref_client.py <server_ip> <own_ip> <my_port> <frames_in_chunk>
"""
import sys

import worker  # the reference module, from baseline/_ref on PYTHONPATH

server_ip, own_ip, my_port, chunk = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
worker.client.my_port = str(my_port)
worker.client.number_of_frames_in_chunk = chunk
w = worker.client(server_ip, own_ip)
print("READY", flush=True)
for line in sys.stdin:
    a = line.strip()
    if not a:
        continue
    if "request" in a:
        w.become_requester(a.split(" ")[1])
    elif a == "end":
        w.stop_requesting_thread()
    else:
        w.exit_threads()
        print("done.", flush=True)
        break
