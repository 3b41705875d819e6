"""Local model-archive resolution (role of hetseq/file_utils.py:78-105 ``cached_path``).

The reference downloads pretrained archives from S3 / HTTP into an etag-keyed cache (boto3 + requests).  This
framework runs on machines without egress, so ``cached_path`` resolves local files, ``file://`` URLs and
previously cached copies only; a remote URL that is not in the cache raises with a clear message instead of
attempting a download.  The cache layout (``<cache_dir>/<sha256(url)>[.<sha256(etag)>]`` + ``.json`` metadata) is
kept so a cache directory populated elsewhere can be dropped in."""
from __future__ import annotations

import json
import os
from hashlib import sha256
from urllib.parse import urlparse

PYTORCH_PRETRAINED_BERT_CACHE = os.environ.get("PYTORCH_PRETRAINED_BERT_CACHE",
                                               os.path.join(os.path.expanduser("~"), ".pytorch_pretrained_bert"))


def url_to_filename(url: str, etag: str = None) -> str:
    name = sha256(url.encode("utf-8")).hexdigest()
    if etag:
        name += "." + sha256(etag.encode("utf-8")).hexdigest()
    return name


def filename_to_url(filename: str, cache_dir: str = None):
    cache_dir = cache_dir or PYTORCH_PRETRAINED_BERT_CACHE
    path = os.path.join(cache_dir, filename)
    if not os.path.exists(path):
        raise EnvironmentError("file {} not found".format(path))
    meta = path + ".json"
    if not os.path.exists(meta):
        raise EnvironmentError("file {} not found".format(meta))
    with open(meta, encoding="utf-8") as f:
        m = json.load(f)
    return m["url"], m.get("etag")


def cached_path(url_or_filename, cache_dir: str = None) -> str:
    cache_dir = cache_dir or PYTORCH_PRETRAINED_BERT_CACHE
    url_or_filename = str(url_or_filename)
    parsed = urlparse(url_or_filename)
    if parsed.scheme in ("", "file"):
        path = parsed.path if parsed.scheme == "file" else url_or_filename
        if os.path.exists(path):
            return path
        raise EnvironmentError("file {} not found".format(path))
    if parsed.scheme in ("http", "https", "s3"):
        prefix = url_to_filename(url_or_filename)
        if os.path.isdir(cache_dir):
            for f in sorted(os.listdir(cache_dir)):
                if f.startswith(prefix) and not f.endswith(".json"):
                    return os.path.join(cache_dir, f)
        raise EnvironmentError("{} is not in the local cache {} and this build never downloads (no network); copy the "
                               "archive next to your config or pre-populate the cache".format(url_or_filename, cache_dir))
    raise ValueError("unable to parse {} as a URL or as a local path".format(url_or_filename))
