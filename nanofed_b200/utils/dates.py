"""Clock helper (reference ``nanofed/utils/dates.py:4-5``)."""

from datetime import datetime, timezone


def get_current_time() -> datetime:
    """Timezone-aware UTC ``now`` used for every timestamp the framework writes."""
    return datetime.now(timezone.utc)
