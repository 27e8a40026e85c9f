// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.cpp).
// Thin C entry points over the reference's OWN codec sources, compiled from where
// they lie under /root/reference/src by oracle/Makefile (never copied here):
//   bit_stream.h (BitStream::push), prefix_code.cpp (createCanonicalPrefixCodes),
//   huffman_table_decoder.h + bit_stream_reader.h (HuffmanTableDecoder).
// Used to validate the restatement in oracle.cpp and to generate tests/golden/.
#include <stddef.h>
#include <algorithm>
#include <limits>

#include "bit_stream.h"
#include "huffman_table_decoder.h"
#include "prefix_code.h"

#include <cstring>

extern "C" {

// BitStream (src/bit_stream.h:12-43): push n (code, bitsCount) pairs, return bytes.
uint64_t ref_bitstream_pack(const uint16_t* codes, const uint8_t* lens, uint64_t n, uint8_t* out, uint64_t cap)
{
    memb::BitStream stream;
    for (uint64_t i = 0; i < n; ++i) {
        stream.push(memb::PrefixCode{codes[i], lens[i]});
    }
    const std::vector<uint8_t> data = stream.data();
    std::memcpy(out, data.data(), std::min<uint64_t>(cap, data.size()));
    return data.size();
}

// createCanonicalPrefixCodes (src/prefix_code.cpp:5-22) with lengths derived from
// size_offsets exactly as HuffmanTableDecoder's ctor does (huffman_table_decoder.h:31-37).
void ref_canonical_codes(const uint8_t* keys, uint32_t nkeys, const uint32_t* size_offsets, uint32_t nsizes,
                         uint16_t* codes, uint8_t* lens)
{
    (void)nsizes;
    std::vector<memb::CodeInfo> lengths;
    size_t current = 0;
    for (uint32_t i = 0; i < nkeys; ++i) {
        while (i >= size_offsets[current]) {
            ++current;
        }
        lengths.push_back({keys[i], current});
    }
    const auto book = memb::createCanonicalPrefixCodes(lengths);
    for (uint32_t i = 0; i < nkeys; ++i) {
        codes[i] = book.at(keys[i]).code;
        lens[i] = static_cast<uint8_t>(book.at(keys[i]).bitsCount);
    }
}

// HuffmanTableDecoder::decode/next (src/huffman_table_decoder.h:96-118)
void ref_decode_stream(const uint8_t* keys, uint32_t nkeys, const uint32_t* size_offsets, uint32_t nsizes,
                       int table_bits, const uint8_t* data, uint64_t size, uint64_t count, uint8_t* out)
{
    const std::vector<uint8_t> k(keys, keys + nkeys);
    const std::vector<uint32_t> so(size_offsets, size_offsets + nsizes);
    memb::HuffmanTableDecoder decoder(k, so, static_cast<size_t>(table_bits));
    auto state = decoder.decode(data, size);
    for (uint64_t i = 0; i < count; ++i) {
        out[i] = decoder.next(state);
    }
}

}  // extern "C"
