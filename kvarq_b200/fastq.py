"""Host-side FastQ helpers mirroring the reference's ``kvarq.fastq`` names the engine needs.

Only the exception type lives here for now (csrc/workhorse.c:1598-1600 imports it from
``kvarq.fastq``); variant sniffing (kvarq/fastq.py:176-238) is a pre-flight step outside the
scanning path.
"""


class FastqFileFormatException(Exception):
    """Raised for records not starting with '@' / third lines not starting with '+'
    (kvarq/fastq.py, raised from csrc/workhorse.c:1037-1048)."""
