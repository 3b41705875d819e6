"""Import stub: the reference imports boto3 at module import (hetseq/file_utils.py:13) but only uses it for S3
downloads, which cannot happen here (no network)."""


def resource(*a, **k):
    raise RuntimeError("boto3 is stubbed: no network on this box")


def client(*a, **k):
    raise RuntimeError("boto3 is stubbed: no network on this box")
