"""chiaswarm_b200 -- a Blackwell (B200, sm_100a) native diffusion-serving worker with the job API of
chiaSWARM (reference: swarm/__init__.py:1 carries the wire version sent to the hive)."""

# Version reported to the hive (query param, user-agent, every result). Kept equal to the
# reference's so existing hives accept the worker (swarm/__init__.py:1).
__version__ = "0.23.6"
# Version of this framework itself.
__framework_version__ = "0.1.0"
