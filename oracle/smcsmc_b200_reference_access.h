/* TEST INFRASTRUCTURE ONLY.  Included by the patched reference driver (tools/patch_reference_driver.py) after the
 * reference's own headers.  Segment::segment_state() is private with ParticleContainer as its friend; a maintainer applying
 * the patch would make it public.  The compile check reaches it without touching the reference's headers, through the
 * explicit-instantiation rule (access checks do not apply to the arguments of an explicit template instantiation). */
#ifndef SMCSMC_B200_REFERENCE_ACCESS_H
#define SMCSMC_B200_REFERENCE_ACCESS_H
namespace b200_access {
template <typename Tag, typename Tag::type M>
struct Open { friend typename Tag::type get(Tag) { return M; } };
struct SegmentState { typedef Segment_State (Segment::*type)() const; friend type get(SegmentState); };
template struct Open<SegmentState, &Segment::segment_state>;
}  // namespace b200_access
static inline bool b200_segment_is_site(const Segment* s) {
  return (s->*get(b200_access::SegmentState()))() == SEGMENT_INVARIANT;
}
#endif
