"""IPlotter interface (reference: veles/iplotter.py)."""
from .verified import Interface


class IPlotter(Interface):
    def redraw(self):
        """Render the filled data."""

    def fill(self):
        """Copy the data to draw."""
