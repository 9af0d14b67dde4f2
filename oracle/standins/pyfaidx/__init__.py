"""TEST INFRASTRUCTURE ONLY -- stand-in for the absent third-party package `pyfaidx`
(reference call sites: /root/reference/src/utils/Pileup.py:41,415).

Fasta.get_seq(name, start, end) is 1-based inclusive; with as_raw=True it returns a plain str and with
sequence_always_upper=True it upper-cases.  Never imported by the product package.
"""


def register_memory_fasta(path, seqs):
    """Make Fasta(path) serve the dict name -> str."""
    Fasta._cache[path] = dict(seqs)


class Fasta(object):
    _cache = {}

    def __init__(self, path, as_raw=False, sequence_always_upper=False, **kw):
        if path not in Fasta._cache:
            seqs, name, chunks = {}, None, []
            with open(path) as fh:
                for line in fh:
                    if line.startswith(">"):
                        if name is not None:
                            seqs[name] = "".join(chunks)
                        name, chunks = line[1:].split()[0], []
                    else:
                        chunks.append(line.strip())
            if name is not None:
                seqs[name] = "".join(chunks)
            Fasta._cache[path] = seqs
        self._seqs = Fasta._cache[path]
        self._upper = sequence_always_upper

    def get_seq(self, name, start, end, rc=False):
        s = self._seqs[name][start - 1:end]
        return s.upper() if self._upper else s
