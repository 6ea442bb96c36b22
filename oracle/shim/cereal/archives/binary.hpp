// TEST INFRASTRUCTURE ONLY: stub of cereal (only named by the reference's dead FM-index classes)
#pragma once
#include <ostream>
namespace cereal {
class BinaryOutputArchive {
public:
    explicit BinaryOutputArchive(std::ostream&) {}
    template <class... T>
    void operator()(T const&...) {}
};
}
