"""Machine-readable metric sink with the SummaryWriter ``add_scalar`` / ``add_text`` API."""
import json
import time


class JsonlScalarWriter:
    def __init__(self, path):
        self._f = open(path, "a")

    def add_scalar(self, name, value, step=None):
        try:
            value = float(value)
        except (TypeError, ValueError):
            return
        self._f.write(json.dumps({"t": time.time(), "name": name, "value": value, "step": step}) + "\n")

    def add_text(self, name, text, global_step=None):
        self._f.write(json.dumps({"t": time.time(), "name": name, "text": str(text), "step": global_step}) + "\n")

    def flush(self):
        self._f.flush()

    def flush_all(self):
        self._f.flush()

    def close(self):
        self._f.close()
