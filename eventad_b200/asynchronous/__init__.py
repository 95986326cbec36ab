"""Asynchronous (incremental) DAGR layers, SURVEY 8 f-1 (/root/reference/src/dagr/asynchronous)."""
