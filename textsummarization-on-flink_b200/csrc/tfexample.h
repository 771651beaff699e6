// tf.Example wire codec for the rows that cross the operator boundary.
//
// In the reference every row travels as a serialized tf.Example: the input queue carries
// {uuid, article, reference} bytes features (batcher.py:554-557, parsed with FixedLenFeature([], tf.string)) and
// FlinkWriter.write_result emits {uuid, article, summary, reference} (flink_writer.py:26-34); the Java side
// configures the same coding through CodingUtils.configureExampleCoding (util/CodingUtils.java:131-145, 196-206).
// This is the protobuf subset those messages use, dependency-free:
//   Example  { Features features = 1; }
//   Features { map<string, Feature> feature = 1; }      (map entry: key = 1, value = 2)
//   Feature  { oneof kind { BytesList bytes_list = 1; FloatList float_list = 2; Int64List int64_list = 3; } }
//   BytesList { repeated bytes value = 1; }
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace pgn {

struct BytesFeature { std::string_view name; std::string_view value; };

// Collects every feature that is a BytesList with exactly one value (what FixedLenFeature([], tf.string)
// accepts); other kinds and unknown fields are skipped.  Views point into `buf`.  Returns "" or an error.
std::string parse_example(const uint8_t* buf, size_t len, std::vector<BytesFeature>* out);
// Serializes the features in the given order (one BytesList value each).
void write_example(const std::vector<BytesFeature>& feats, std::string* out);

}  // namespace pgn
