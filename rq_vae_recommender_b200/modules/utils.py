"""`eval_mode` decorator used by the tokenizer (reference: modules/utils.py:31-39)."""
import functools


def eval_mode(fn):
    @functools.wraps(fn)
    def inner(self, *args, **kwargs):
        was_training = self.training
        self.eval()
        try:
            return fn(self, *args, **kwargs)
        finally:
            self.train(was_training)
    return inner
