"""Privacy error hierarchy (reference ``privacy/exceptions.py:1-22``)."""


class PrivacyError(Exception):
    """Root of privacy-related failures."""


class PrivacyBudgetExceededError(PrivacyError):
    """The (epsilon, delta) budget has been used up."""


class PrivacyConfigurationError(PrivacyError):
    """A privacy configuration is inconsistent."""


class NoiseGenerationError(PrivacyError):
    """A noise sample could not be drawn."""
