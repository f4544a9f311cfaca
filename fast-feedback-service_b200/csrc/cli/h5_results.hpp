// h5_results.hpp -- the reflection table the reference writes with --save-h5
// (spotfinder/spotfinder.cc:1219-1305 through dx2 ReflectionTable, dx2/dx2/reflection.cxx:14-21,
// 117-131, 181-260 and dx2/dx2/h5/h5write.cxx:60-100), on h5lite instead of libhdf5.
//
// Same file shape: group `dials/processing/group_0` with attributes `experiment_ids`
// (uint64[n]) and `identifiers` (variable-length UTF-8 strings[n]), one contiguous dataset per
// column; an N x 1 column is stored with shape (N,) (reflection.cxx:236-243).  Column types are
// the ones dx2 registers (h5dispatch.hpp:64-77): int, int64_t, uint64_t, double.
#pragma once
#include <array>
#include <iomanip>
#include <random>
#include <sstream>
#include <typeindex>

#include "h5lite.hpp"

namespace h5results {

// dx2/dx2/utils.cxx:44-68: 16 random bytes, printed most significant first as 8-4-4-4-12
inline std::string ersatz_uuid4() {
    std::array<unsigned char, 16> bytes;
    std::random_device rd;
    std::uniform_int_distribution<int> dist(0, 255);
    for (auto& b : bytes) b = (unsigned char)dist(rd);
    std::ostringstream oss;
    for (auto it = bytes.rbegin(); it != bytes.rend(); ++it) oss << std::hex << std::setw(2) << std::setfill('0') << (int)*it;
    const std::string hex = oss.str();
    return hex.substr(0, 8) + "-" + hex.substr(8, 4) + "-" + hex.substr(12, 4) + "-" + hex.substr(16, 4) + "-" + hex.substr(20, 12);
}

class Table {
    struct Column {
        std::string name;
        h5lite::Datatype dtype;
        std::vector<uint64_t> shape;
        std::vector<uint8_t> bytes;
    };
    std::vector<Column> columns_;
    std::vector<uint64_t> experiment_ids_;
    std::vector<std::string> identifiers_;
    uint64_t max_experiment_id_ = 0;

    template <typename T>
    static h5lite::Datatype type_of() {
        if (std::is_same<T, double>::value) return h5lite::Datatype::f64();
        if (std::is_same<T, int>::value) return h5lite::Datatype::i32();
        if (std::is_same<T, int64_t>::value) return h5lite::Datatype::i64();
        if (std::is_same<T, uint64_t>::value) return h5lite::Datatype::u64();
        throw h5lite::Error("Attempted to add column with unsupported type");
    }

  public:
    // A fresh table owns one experiment (reflection.cxx:14-17)
    Table() { generate_new_attributes(); }

    std::pair<uint64_t, std::string> generate_new_attributes() {
        const uint64_t id = max_experiment_id_++;
        std::string uuid = ersatz_uuid4();
        experiment_ids_.push_back(id);
        identifiers_.push_back(uuid);
        return {id, uuid};
    }
    const std::vector<uint64_t>& get_experiment_ids() const { return experiment_ids_; }
    const std::vector<std::string>& get_identifiers() const { return identifiers_; }

    template <typename T>
    void add_column(const std::string& name, size_t rows, size_t cols, const std::vector<T>& data) {
        if (data.size() != rows * cols) throw h5lite::Error("Column '" + name + "': data size does not match its shape");
        if (!columns_.empty() && columns_.front().shape[0] != rows)
            throw h5lite::Error("Column '" + name + "' has a different number of rows");
        Column c;
        c.name = name, c.dtype = type_of<T>(), c.shape = {rows, cols};
        c.bytes.assign((const uint8_t*)data.data(), (const uint8_t*)data.data() + data.size() * sizeof(T));
        columns_.push_back(std::move(c));
    }

    void write(const std::string& filename, const std::string& group = "/dials/processing/group_0") const {
        if (experiment_ids_.empty() || identifiers_.empty())  // h5write.cxx:64-67
            throw h5lite::Error("Experiment IDs and identifiers must not be empty.");
        h5lite::Writer w;
        auto* g = w.group(group);
        h5lite::Writer::attr_raw(g, "experiment_ids", h5lite::Datatype::u64(), {experiment_ids_.size()}, experiment_ids_.data());
        h5lite::Writer::attr_strings(g, "identifiers", identifiers_);
        for (const auto& c : columns_) {
            std::vector<uint64_t> shape = c.shape;
            if (shape.size() == 2 && shape[1] == 1) shape = {shape[0]};
            w.dataset(g, c.name, c.dtype, shape, c.bytes.data());
        }
        w.save(filename);
    }
};

}  // namespace h5results
