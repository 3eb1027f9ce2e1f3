"""Minimal DataLog with the reference's interface (mjrl/utils/logger.py:10-57): dict of lists,
`log_kv`, `save_log` (pickle + csv), `get_current_log`, `read_log`.  No matplotlib."""
import csv
import pickle


class DataLog:
    def __init__(self):
        self.log = {}
        self.max_len = 0

    def log_kv(self, key, value):
        if key not in self.log:
            self.log[key] = []
        self.log[key].append(value)
        if len(self.log[key]) > self.max_len:
            self.max_len = self.max_len + 1

    def save_log(self, save_path):
        pickle.dump(self.log, open(save_path + "/log.pickle", "wb"))
        with open(save_path + "/log.csv", "w") as f:
            w = csv.DictWriter(f, fieldnames=list(self.log.keys()))
            w.writeheader()
            for row in range(self.max_len):
                w.writerow({k: self.log[k][row] for k in self.log if row < len(self.log[k])})

    def get_current_log(self):
        return {k: v[-1] for k, v in self.log.items()}

    def read_log(self, log_path):
        with open(log_path) as f:
            rows = list(csv.DictReader(f))
        data = {k: [] for k in (rows[0].keys() if rows else [])}
        for r in rows:
            for k in data:
                try:
                    data[k].append(float(r[k]))
                except (TypeError, ValueError):
                    pass
        self.log = data
        self.max_len = len(rows)
