// Driver for intercession_b200/host/jump_history.hpp: reads a script of jump events from stdin, prints one
// result per line.  The Python test feeds the same script to the reference-side harness (oracle/_ref).
//   D entity trip x y z   cache_departure           -> "D"
//   A trip x y z          on_arrival                -> "A <paradox 0/1>"
//   R entity x y z        on_jump_request           -> "R <kind 0..3> <tripId>"
//   C                     clear_departures          -> "C"
#include "../../intercession_b200/host/jump_history.hpp"

#include <cstdio>

int main()
{
    icx_host::JumpHistory jh;
    char op;
    while (std::scanf(" %c", &op) == 1) {
        unsigned entity = 0, trip = 0;
        unsigned bits[3];
        float o[3];
        auto origin = [&]() {       // origins travel as raw FP32 bit patterns: no decimal round trip
            if (std::scanf("%x %x %x", &bits[0], &bits[1], &bits[2]) != 3) return false;
            for (int k = 0; k < 3; k++) __builtin_memcpy(&o[k], &bits[k], 4);
            return true;
        };
        if (op == 'D') {
            if (std::scanf("%u %u", &entity, &trip) != 2 || !origin()) return 1;
            jh.cache_departure(entity, trip, o);
            std::puts("D");
        } else if (op == 'A') {
            if (std::scanf("%u", &trip) != 1 || !origin()) return 1;
            std::printf("A %d\n", jh.on_arrival(trip, o) ? 1 : 0);
        } else if (op == 'R') {
            if (std::scanf("%u", &entity) != 1 || !origin()) return 1;
            uint32_t t = 0;
            const int kind = (int)jh.on_jump_request(entity, o, &t);
            std::printf("R %d %u\n", kind, t);
        } else if (op == 'C') {
            jh.clear_departures();
            std::puts("C");
        } else return 1;
    }
    return 0;
}
