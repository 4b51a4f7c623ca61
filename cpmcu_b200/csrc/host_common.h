// Host-side arithmetic shared by the dispatcher and the sizing helpers.
#pragma once

namespace b200 {

// Upper bound on the number of CTAs whose unit ranges intersect one column of `kblks` units when
// `units` units are split evenly over `grid` CTAs (every CTA gets floor or ceil of units/grid):
// one partial run at each end plus the whole runs between them.  This is the per-column slot
// stride reserved in c_tmp; the kernel computes the exact count per column itself.
inline int decode_slot_max(int kblks, long long units, int grid) {
  long long minrun = units / grid;
  if (minrun < 1) minrun = 1;
  return static_cast<int>((kblks - 1 + minrun - 1) / minrun) + 1;
}

}  // namespace b200
