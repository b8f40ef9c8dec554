"""B200-native distributed video inference (DistributedVolunteerComputing capabilities)."""
import os as _os

# One hardware queue per CUDA stream (the default is 8): the mailbox data plane keeps one stream per (worker, lane) and
# parks bounded wait kernels on them; with aliased queues a parked wait would hold up an unrelated lane until its
# time-out.  Read by the driver when the CUDA context is created, so it has to be in the environment before that.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
