"""Units that contribute to the ``--result-file`` JSON (reference: veles/result_provider.py:41-58)."""
from .verified import Interface


class IResultProvider(Interface):
    def get_metric_names(self):
        """Set of metric names this unit reports."""

    def get_metric_values(self):
        """Dict name -> JSON-serialisable value."""
