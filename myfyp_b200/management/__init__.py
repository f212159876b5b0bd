"""Management / observability layer: logger singleton, metric stores, monitors."""
