// compat/pairings.h - `#include <pairings.h>` of the reference (walks/pairings.h:7-128) over the host scalar forms of
// the library's pairing ops (wmh_szudzik_pair / wmh_szudzik_unpair: same arithmetic as the device ops, exact integer
// square root).  Results are reduced to Type exactly as the reference's arithmetic in Type wraps.
#ifndef WHARFMH_B200_COMPAT_PAIRINGS_H
#define WHARFMH_B200_COMPAT_PAIRINGS_H

#include "../wharfmh_b200.h"

#include <cstdint>
#include <tuple>
#include <utility>

namespace pairings
{
    template <class Type>
    class Szudzik
    {
    public:
        static Type pair(const std::pair<Type, Type>& pair)                                              // pairings.h:27-79
        { return static_cast<Type>(wmh_szudzik_pair(static_cast<uint64_t>(pair.first), static_cast<uint64_t>(pair.second))); }

        static Type pair_triplet(const std::tuple<Type, Type, Type>& tuple)                              // :88-91
        { return pair({pair({std::get<0>(tuple), std::get<1>(tuple)}), std::get<2>(tuple)}); }

        static std::pair<Type, Type> unpair(const Type& encoded_value)                                   // :100-113
        {
            uint64_t x = 0, y = 0;
            wmh_szudzik_unpair(static_cast<uint64_t>(encoded_value), &x, &y);
            return {static_cast<Type>(x), static_cast<Type>(y)};
        }

        static std::tuple<Type, Type, Type> unpair_triplet(const Type& encoded_value)                    // :122-128
        {
            auto first_pair = unpair(encoded_value);
            auto second_pair = unpair(first_pair.first);
            return std::make_tuple(second_pair.first, second_pair.second, first_pair.second);
        }
    };
}

#endif
