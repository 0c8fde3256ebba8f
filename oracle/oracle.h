/* oracle/oracle.h — C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under cuda_ray_tracer_b200/ may include, link or call this.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg use it, as the checker.
 *
 * The oracle is a sequential CPU restatement of the reference's hot path
 * (GJ0407790/cuda_ray_tracer): scene parsing (src/parse.cpp), Morton codes + stable sort
 * (src/lbvh_utils.cu), Karras hierarchy + AABB refit (src/lbvh_builder.cu), BVH traversal
 * (src/bvh_traversal.cu), primitive intersection and camera rays (src/struct.cu), recursive
 * Whitted shading and the two live render kernels (src/draw.cu), colour/RNG helpers
 * (src/helper.cu) and the cuRAND XORWOW generator the reference seeds per pixel.
 * Each function in oracle.cpp cites the reference file:line it follows.
 */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct o_scene o_scene;

/* bounds modes: as shipped the reference never stores the scene bounds (±INF ⇒ all
 * Morton codes 0); "computed" is the bounding box parse.cpp accumulates and then drops. */
enum { O_BOUNDS_REF_SHIPPED = 0, O_BOUNDS_COMPUTED = 1 };
/* sample semantics: the reference's render() as shipped, or true n-spp. */
enum { O_SPP_REF = 0, O_SPP_TRUE = 1 };

/* evaluation order of the three unsequenced standerdD() draws in draw.cu:335-337,395-397:
 * 0 = x,y,z (left to right), 1 = z,y,x (g++ host build of the reference). */
void o_set_args_rtl(int rtl);

o_scene* o_scene_parse(const char* path);            /* NULL on open failure / invalid line */
o_scene* o_scene_new(void);
void     o_scene_free(o_scene*);

/* scalar config access (field order documented in oracle.cpp: o_scene_get_config) */
void o_scene_get_config(const o_scene*, int* ints /*[8]*/, float* floats /*[32]*/);
void o_scene_set_config(o_scene*, const int* ints, const float* floats);
const char* o_scene_filename(const o_scene*);

int o_scene_counts(const o_scene*, int* n_spheres, int* n_triangles, int* n_prims,
                   int* n_suns, int* n_bulbs, int* n_planes);
/* reference-layout host arrays (12-byte vec3, 44-byte Materials).  Pointers stay valid until free. */
const float* o_scene_sphere_c(const o_scene*);   /* [ns][3] */
const float* o_scene_sphere_r(const o_scene*);   /* [ns]    */
const float* o_scene_sphere_mat(const o_scene*); /* [ns][11] */
const float* o_scene_tri(const o_scene*, int which /*0..5 = p0,p1,p2,nor,e1,e2*/); /* [nt][3] */
const float* o_scene_tri_mat(const o_scene*);    /* [nt][11] */
const int32_t* o_scene_refs(const o_scene*);     /* [n][2] = (type, id_in_type_array), file order */
const float* o_scene_suns(const o_scene*);       /* [n][6] dir, color */
const float* o_scene_bulbs(const o_scene*);      /* [n][6] point, color */
const float* o_scene_planes(const o_scene*);     /* [n][21] a,b,c,d, nor3, point3, mat11 */
void o_scene_bounds(const o_scene*, int mode, float* mn3, float* mx3);

/* programmatic scene construction (synthetic soups / sphere fields) */
void o_scene_add_triangles(o_scene*, int n, const float* p0, const float* p1, const float* p2,
                           const float* mat11 /*[n][11]*/);
void o_scene_add_spheres(o_scene*, int n, const float* c, const float* r, const float* mat11);
void o_scene_add_sun(o_scene*, const float* dir3, const float* color3);
void o_scene_add_bulb(o_scene*, const float* p3, const float* color3);
void o_scene_add_plane(o_scene*, const float* abcd, const float* mat11);

/* Triangle constructor arithmetic (include/object.cuh:177-191) for n triangles. */
void o_triangle_setup(int n, const float* p0, const float* p1, const float* p2,
                      float* nor, float* e1, float* e2);

/* ---- LBVH build (K1..K6) ---- */
int o_build(o_scene*, int bounds_mode, int morton_bits /*30 or 63*/);
int o_num_nodes(const o_scene*);
const uint32_t* o_codes30(const o_scene*);        /* sorted, [n] (morton_bits==30) */
const uint64_t* o_codes63(const o_scene*);        /* sorted, [n] (morton_bits==63) */
const uint32_t* o_perm(const o_scene*);           /* sorted position -> file-order primitive index */
const int32_t*  o_sorted_refs(const o_scene*);    /* [n][2] */
const float*    o_node_bbox(const o_scene*);      /* [2n-1][6] xmin,xmax,ymin,ymax,zmin,zmax */
const uint32_t* o_node_left(const o_scene*);      /* [n-1] */
const uint32_t* o_node_right(const o_scene*);     /* [n-1] */
const int32_t*  o_node_parent(const o_scene*);    /* [2n-1], root = -1 */
/* unsorted codes of file-order primitives (K1 alone) */
void o_morton_unsorted(const o_scene*, int bounds_mode, int morton_bits, uint64_t* out);

/* ---- XORWOW (cuRAND) ---- */
void o_xorwow_init(unsigned long long seed, unsigned long long subsequence, uint32_t* state6 /*v0..v4,d*/);
void o_xorwow_draws(uint32_t* state6, int n, uint32_t* out);

/* ---- trace / render ---- */
typedef struct {
  unsigned long long rays;          /* hitNearest calls with bounce != 0 (every BVH traversal) */
  unsigned long long nodes_visited; /* loop iterations of traverse_lbvh (leaf or internal) */
  unsigned long long leaf_tests;    /* primitive intersection tests */
  unsigned long long shaded_hits;   /* hits whose material is consumed by shading */
  int max_stack;
} o_counters;

/* Closest hit for explicit rays: id = file-order primitive index, -1 miss, -2-p plane p. */
void o_trace(const o_scene*, int n, const float* org3, const float* dir3 /*normalised by caller*/,
             int32_t* hit_id, float* hit_t, o_counters*);

/* Full render. rgba [h][w][4] u8, hit_id [h][w] (primary ray of sample 0), written[h][w] u8 flag:
 * 1 where the reference's launch shape writes the pixel (F3).  w/h/aa <0 ⇒ scene's own values.
 * y0,y1 restrict to rows [y0,y1) (others untouched); threads = host threads to use. */
int o_render(const o_scene*, int w, int h, int aa, int spp_mode, int y0, int y1, int threads,
             uint8_t* rgba, int32_t* hit_id, uint8_t* written, float* linear_rgba /*optional [h][w][4]*/,
             o_counters*);

#ifdef __cplusplus
}
#endif
#endif
