"""`CodecMixin.get_delay()` is called at modded_dac.py:859; its value is unused on decode."""


class CodecMixin:
    def get_delay(self):
        return 0
