"""The JSON edge either side of the stage in the reference's demo application -- SURVEY.md section 8f row 4.

``App.startInference`` (src/main/java/me/littlebo/App.java:108-131) reads Kafka records with
``MessageDeserializationSchema`` (MessageDeserializationSchema.java:22-32), runs ``TFModel.transform`` on the columns
``uuid, article, reference`` and writes the result with ``MessageSerializationSchema``
(MessageSerializationSchema.java:14-27).  Both schemas map a JSON object with the four string properties of
``Message`` (Message.java:11-19) to and from a 4-field row.  Kafka itself is out of scope; this module is the serde and
the dataflow around the stage, so that a stream of the demo's JSON records can be pushed through the library.

Jackson specifics that are reproduced: property order ``uuid, article, summary, reference`` (declaration order), no
whitespace, ``null`` for an absent value, non-ASCII characters written as UTF-8 (not escaped), control characters as
``\\b \\t \\n \\f \\r`` or ``\\u00XX`` with upper-case hex digits, an unknown property is an error
(``FAIL_ON_UNKNOWN_PROPERTIES``), a missing property reads as ``None``.  PARITY UNPINNED: no JVM here, so this is
checked against the classes' source and Jackson's documented defaults, not against their output.
"""
from __future__ import annotations

import json
from typing import Iterable, Iterator, List, Optional, Sequence, Tuple

FIELDS = ("uuid", "article", "summary", "reference")          # Message.java:11-18
Row = Tuple[Optional[str], Optional[str], Optional[str], Optional[str]]

_SHORT = {0x08: "\\b", 0x09: "\\t", 0x0A: "\\n", 0x0C: "\\f", 0x0D: "\\r", 0x22: '\\"', 0x5C: "\\\\"}


def _json_string(s: Optional[str]) -> str:
    if s is None:
        return "null"
    out = ['"']
    for ch in s:
        c = ord(ch)
        if c in _SHORT:
            out.append(_SHORT[c])
        elif c < 0x20:
            out.append("\\u%04X" % c)
        else:
            out.append(ch)
    out.append('"')
    return "".join(out)


def serialize(row: Sequence[Optional[str]]) -> bytes:
    """MessageSerializationSchema.serialize (:14-27): row -> JSON bytes; ``b""`` when the row cannot be written
    (the reference logs and returns an empty array)."""
    try:
        if len(row) != 4:
            raise ValueError("row must have 4 fields")
        vals = [v.decode("utf-8") if isinstance(v, (bytes, bytearray)) else v for v in row]
        if any(v is not None and not isinstance(v, str) for v in vals):
            raise TypeError("fields must be strings")
        return ("{" + ",".join('"%s":%s' % (k, _json_string(v)) for k, v in zip(FIELDS, vals)) + "}").encode("utf-8")
    except (ValueError, TypeError, UnicodeDecodeError):
        return b""


class MessageDeserializer:
    """MessageDeserializationSchema (:13-41): JSON bytes -> row, counting records; ``is_end_of_stream`` turns true once
    MORE than ``max_count`` records have been read (the reference tests ``counter > maxCount``, :35-38)."""

    def __init__(self, max_count: int):
        self.max_count = int(max_count)
        self.counter = 0

    def deserialize(self, data: bytes) -> Row:
        obj = json.loads(data.decode("utf-8"))
        if not isinstance(obj, dict):
            raise ValueError("a Message is a JSON object")
        unknown = [k for k in obj if k not in FIELDS]
        if unknown:
            raise ValueError("Unrecognized field \"%s\" (class me.littlebo.Message)" % unknown[0])
        vals: List[Optional[str]] = []
        for k in FIELDS:
            v = obj.get(k)
            if isinstance(v, (dict, list)):
                raise ValueError("Cannot deserialize a String from an object or array (property \"%s\")" % k)
            if v is not None and not isinstance(v, str):                 # Jackson coerces scalars to their text
                v = str(v).lower() if isinstance(v, bool) else str(v)
            vals.append(v)
        self.counter += 1
        return tuple(vals)                                       # type: ignore[return-value]

    def is_end_of_stream(self) -> bool:
        return self.counter > self.max_count


def transform_messages(model, records: Iterable[bytes], max_count: int = 8, devices=None) -> Iterator[bytes]:
    """App.startInference's dataflow without Kafka (:116-130): JSON records -> rows ``uuid, article, summary,
    reference`` -> ``select uuid, article, reference`` -> ``model.transform`` -> JSON records.  ``model`` is a
    ``pipeline.TFModel``; the stream ends like the reference's source, after more than ``max_count`` records."""
    de = MessageDeserializer(max_count)

    def rows():
        for rec in records:
            uuid, article, _summary, reference = de.deserialize(rec)
            if de.is_end_of_stream():                             # the record that trips the limit is not emitted
                return
            yield {"uuid": uuid or "", "article": article or "", "reference": reference or ""}

    for out in model.transform(rows(), devices=devices):
        yield serialize([out.get(k) for k in FIELDS])
