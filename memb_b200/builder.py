"""Builder — mirrors python/memb/builder.py:4-39 (write side; fixture generator)."""
from . import _memb


class Builder:
    '''Quantizes embeddings with the given precision and saves a memb file.

    Parameters
    ----------
    dim : int
        Dimension of word vectors
    storage_type : str
        'full', 'uniform' or 'trained'
    bits_per_weight : int
        Bits per weight; values beyond the range a strategy accepts are clamped
        by the strategy (levels = min(1 << bits, 255))
    device : int or None, keyword only, additive
        CUDA ordinal for the GPU converter: the bulk passes of the 'trained' compressor
        (predict, Huffman-encode every row) run on that GPU at save(); None = host threads.
        The file is byte-identical either way.
    '''

    def __init__(self, dim, storage_type='trained', bits_per_weight=4, *, device=None):
        self._impl = _memb.Builder(dim, storage_type, bits_per_weight, -1 if device is None else int(device))

    def add_word(self, word, vector):
        '''Add one word; vector is a 1-D float32 array of length dim.'''
        self._impl.add_word(word, vector)

    def add_words(self, words, matrix):
        '''Additive bulk form: list of str + (len(words), dim) float32 matrix.'''
        self._impl.add_words(list(words), matrix)

    def save(self, filename):
        '''Compress the content and write it to `filename` (str or pathlib.Path).'''
        self._impl.save(str(filename))
