"""Rotating file logging (reference: swarm/log_setup.py:5-29 -- 50 MiB x 7 backups, ISO
timestamps). `concurrent_log_handler` is not a dependency here: one writer process owns the file
(rank processes log through the parent), so stdlib RotatingFileHandler is sufficient. Both
"WARN" (the reference's default string, which it mis-maps to INFO) and "WARNING" are accepted."""
from __future__ import annotations

import logging
from logging.handlers import RotatingFileHandler

_LEVELS = {
    "CRITICAL": logging.CRITICAL,
    "ERROR": logging.ERROR,
    "WARNING": logging.WARNING,
    "WARN": logging.WARNING,
    "INFO": logging.INFO,
    "DEBUG": logging.DEBUG,
}

_FORMAT = "%(asctime)s - %(levelname)s - %(message)s"
_DATEFMT = "%Y-%m-%dT%H:%M:%S"


def setup_logging(log_path, log_level) -> logging.Handler:
    logger = logging.getLogger()
    logger.setLevel(_LEVELS.get(str(log_level).upper(), logging.INFO))
    for h in list(logger.handlers):
        if getattr(h, "_swarm_handler", False):
            logger.removeHandler(h)
            h.close()
    handler = RotatingFileHandler(str(log_path), mode="a", maxBytes=50 * 1024 * 1024, backupCount=7)
    handler.setFormatter(logging.Formatter(fmt=_FORMAT, datefmt=_DATEFMT))
    handler._swarm_handler = True  # type: ignore[attr-defined]
    logger.addHandler(handler)
    return handler
