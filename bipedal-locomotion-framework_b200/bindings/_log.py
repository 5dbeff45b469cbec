import logging

_logger = logging.getLogger("blf")


def log():
    """BipedalLocomotion::log() (src/TextLogging): errors are reported here and the call returns False."""
    return _logger
