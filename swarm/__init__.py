"""Drop-in alias so deployments written against the reference keep working:
`python -m swarm.worker`, `python -m swarm.initialize`, `from swarm.generator import do_work` ...
resolve to the chiaswarm_b200 implementation (reference package: swarm/__init__.py:1)."""
import chiaswarm_b200 as _impl

__version__ = _impl.__version__
__path__ = list(_impl.__path__)
