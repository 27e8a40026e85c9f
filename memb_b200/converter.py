"""Word-vector text -> memb file: the command line of the reference's
tools/converter/memb_converter (flags :39-79, parsing rules :11-33) on top of
memb_b200.Builder, so files can be produced without the reference toolchain.

Text format: one word per line followed by its weights, separated by single spaces
(word2vec / GloVe text).  Lines with fewer than two weights (e.g. the word2vec
"rows dim" header) are skipped, a line whose weights do not parse is reported and
skipped, the dimension is the median weight count and lines of another length are
discarded with a message -- all as in the reference.
"""
import argparse
import sys
from itertools import islice
from statistics import median

import numpy as np

from . import Builder, available_compression_strategies


def read_vectors(filename, max_words=None, out=sys.stdout):
    """[(word, float32 vector)] and the dimension, following the reference's rules."""
    parsed = []
    with open(filename) as source:
        for line in islice(source, max_words):
            fields = line.strip().split(' ')
            if len(fields) <= 2:
                continue
            try:
                parsed.append((fields[0], np.asarray(fields[1:], dtype=np.float32)))
            except Exception as error:  # a weight that is not a number
                print('Exception while parsing: {}'.format(error), file=out)
    if not parsed:
        return [], 0
    dim = int(median(len(vector) for _, vector in parsed))
    kept = [(word, vector) for word, vector in parsed if len(vector) == dim]
    discarded = len(parsed) - len(kept)
    if discarded > 0:
        print('{} items are discarded due to inconsistent vector length'.format(discarded), file=out)
    return kept, dim


def convert(source_filename, dest_filename, quantization, bits_per_weight=4, max_words=None, out=sys.stdout, device=None):
    """Returns the number of words written."""
    vectors, dim = read_vectors(source_filename, max_words, out)
    builder = Builder(dim, quantization, bits_per_weight, device=device)
    written = 0
    for word, vector in vectors:
        try:
            builder.add_word(word, vector)
            written += 1
        except Exception as error:  # duplicate word (src/builder.cpp:12-16 in the reference)
            print('Exception ({}) while trying to add word'.format(error), file=out)
    builder.save(dest_filename)
    return written


def parse_args(argv=None):
    parser = argparse.ArgumentParser(description='Convert word vectors to quantized binary format')
    parser.add_argument('--from', dest='source_filename', required=True, help='Source filename')
    parser.add_argument('--to', dest='dest_filename', required=True, help='Destination filename')
    parser.add_argument('--quantization', dest='quantization', required=True,
                        choices=available_compression_strategies(),
                        help='Quantization strategy; "trained" gives the smallest files at the best quality')
    parser.add_argument('--bits-per-weight', dest='bits_per_weight', type=int, default=4,
                        help='Bits per weight; a value outside the range of the strategy is clamped to it')
    parser.add_argument('--max-words', dest='max_words', type=int,
                        help='Convert at most this many lines (default: all)')
    parser.add_argument('--device', dest='device', type=int, default=None,
                        help='(additive) CUDA ordinal: encode the "trained" storage on that GPU; the file is the same')
    return parser.parse_args(argv)


def main(argv=None):
    args = parse_args(argv)
    convert(args.source_filename, args.dest_filename, args.quantization, args.bits_per_weight, args.max_words,
            device=args.device)
    return 0


if __name__ == '__main__':
    sys.exit(main())
