class Logger(object):
    def log_metrics(self, metrics, step=None):
        pass

    def log_hyperparams(self, params):
        pass

    def save(self):
        pass

    def finalize(self, status):
        pass


def rank_zero_experiment(fn):
    return fn
