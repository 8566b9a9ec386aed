// PrestoPage wire format for the vectors this path produces: what PartitionedOutput hands to an
// Exchange when a GPU partial aggregation feeds a CPU Velox (or Presto) stage, and what a GPU final aggregation's host
// side reads when its input comes from one.  Restates the layout of velox/serializers/PrestoSerializer.cpp:
//   page   (:1947-1949, :1964-2030)  numRows(4) | codec(1) | uncompressedSize(4) | size(4) | checksum(8) | data
//   data                              numColumns(4) | column ...
//   column (:1018-1023, :1070-1147)  nameLength(4) | encoding name | body
//     fixed width ("BYTE_ARRAY" bool / tinyint, "SHORT_ARRAY", "INT_ARRAY" int / real / date, "LONG_ARRAY" bigint / double,
//                  :102-137):         positions(4) | nulls | values of the NON-NULL rows
//     "VARIABLE_WIDTH" (:1133-1140):  positions(4) | end offsets (4 each; a NULL row repeats the previous one) | nulls |
//                                     totalBytes(4) | bytes
//     "ROW" (:1097-1112, :1362-1391): numChildren(4) | child columns over the NON-NULL rows | positions(4) |
//                                     offsets (positions + 1, first 0, a NULL row repeats) | nulls
//   nulls  (:1149-1158, :943)        hasNulls(1) [| one bit per row, 1 = NULL, most significant bit first]
//   checksum (:34-45, :84-88)        codec bit 4 set: CRC-32 over data, then the codec byte, numRows, uncompressedSize
// Compression and encryption (codec bits 1, 2) are not produced and are rejected on input.  Dictionary / constant
// wrappers are flattened on output (always legal: "RLE" / "DICTIONARY" encodings are optional) and rejected on input.
#pragma once
#include <string>

#include "Vectors.h"

namespace vb200 {
namespace serializer {
namespace presto {

// One page holding every row of `rows`.
std::string serializePage(const RowVectorPtr& rows, bool withChecksum = false);
// The page back as flat vectors of `type` (throws VeloxRuntimeError on a malformed page or a checksum mismatch).
RowVectorPtr deserializePage(const std::string& page, const RowTypePtr& type);

// PartitionedOutput for one BIGINT key column (velox/exec/PartitionedOutput.cpp + HashPartitionFunction.cpp:59-92): row r
// goes to partition hash(key[r]) % numPartitions with the reference's own hash (vgpu_partition_of), each partition's
// rows become one page (an empty string for a partition without rows).
std::vector<std::string> partitionAndSerialize(const RowVectorPtr& rows, int keyChannel, int numPartitions, bool withChecksum = false);

uint32_t crc32(const void* data, size_t n, uint32_t seed = 0); // the standard reflected CRC-32 (boost::crc_32_type)

} // namespace presto
} // namespace serializer
} // namespace vb200
