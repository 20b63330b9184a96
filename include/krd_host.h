/* krd_host.h — C ABI of the host-side pieces around the KR distance call (no CUDA):
 * jplace ingest -> flat mass arrays, and the distance-matrix text writer.
 *
 * Replaces, for `gappa analyze krd`:
 *   JplaceInputOptions::mass_tree_set()      gappa src/options/jplace_input.cpp:295-332
 *     (JplaceReader::read                    genesis placement/formats/jplace_reader.cpp:74-115,
 *      convert_sample_to_mass_tree           genesis placement/function/operators.cpp:172-182,
 *      mass_trees_make_average_branch_lengths genesis tree/mass_tree/functions.cpp:271-335)
 *   FileInputOptions::file_paths / base_file_names   gappa src/options/file_input.cpp:88-156
 *   diameter( mass_trees[0] )                genesis tree/common_tree/functions.cpp:178-205
 *   MatrixOutputOptions::write_matrix        gappa src/options/matrix_output.cpp:76-111
 *     (MatrixWriter<double>                  genesis utils/containers/matrix/writer.hpp:102-320)
 *
 * The arrays returned by krd_host_set_copy are exactly the arguments of krd_set_tree /
 * krd_set_masses in krd_b200.h (edges numbered like the reference numbers them).
 */
#ifndef KRD_HOST_H
#define KRD_HOST_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct krd_host_set krd_host_set;

/* Message of the last failed call on this thread (reference wording where one exists). */
const char* krd_host_last_error(void);

/* paths: files and/or directories (directories contribute *.jplace / *.jplace.gz), pooled and
 * sorted.  Returns NULL on error. */
krd_host_set* krd_host_load_jplace(int32_t n_paths, const char* const* paths, int32_t point_mass,
                                   int32_t ignore_multiplicities, int32_t threads);
void krd_host_free_set(krd_host_set* s);

void krd_host_set_sizes(const krd_host_set* s, int32_t* n_edges, int32_t* n_samples, int64_t* n_masses,
                        int32_t* root, double* diameter);
/* Any output pointer may be NULL.  Sizes: parent/child/edge_num/bl [n_edges]; off [n_samples+1];
 * edge/pos/mass [n_masses]; pendant [n_samples]. */
void krd_host_set_copy(const krd_host_set* s, int32_t* parent, int32_t* child, int64_t* edge_num, double* bl,
                       int64_t* off, int32_t* edge, double* pos, double* mass, double* pendant);
const char* krd_host_set_label(const krd_host_set* s, int32_t i);

/* format: "matrix" | "list" | "triangular" (case-insensitive); labels NULL = no labels.
 * Returns the number of uncompressed bytes written, -1 on error. */
int64_t krd_host_write_matrix(const char* path, const double* m, int32_t n, const char* format,
                              const char* const* labels, const char* corner, int32_t gzip, int32_t threads);

/* Secondary surface: placement::earth_movers_distance( SampleSet, p, with_pendant_length )
 * (placement/function/emd.cpp:107-129) = the mass-tree matrix (krd_compute, divisor 1) plus, with
 * with_pendant_length, pendant[i] + pendant[j] in EVERY cell including the diagonal; `pendant` is the
 * per-sample pendant work krd_host_set_copy returns (placement/function/operators.cpp:167). In place. */
void krd_host_add_pendant_work(double* matrix, int32_t n, const double* pendant);

/* The reference's (not correctly rounded) JSON number accumulation, exposed for tests. */
double krd_host_parse_number(const char* text);

#ifdef __cplusplus
}
#endif
#endif /* KRD_HOST_H */
