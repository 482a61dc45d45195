// main of oracle/_ref/ref_path_a: the UNMODIFIED reference driver (token_test.cc compiled with -Dmain=ref_token_test_main, its
// operators, transform steps and code generator) linked with integration/executor_b200.cc instead of executor.cc.
//   ref_path_a <matrix.mtx> <N>                      the reference's own main: test_spmm_thread_total(meta, name, 4, 1) (token_test.cc:1656)
//   ref_path_a <matrix.mtx> <N> <family> <S> <C>     another of the seven driver chains the reference's main keeps commented out
#include "metadata_set.hpp"
#include "config.hpp"

#include <cstdio>
#include <cstdlib>
#include <string>

using namespace std;

int ref_token_test_main(int argc, char **argv);
float test_spmm_thread_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_bitmap(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_thread_bit_map(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_segment(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_block_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_block_bitmap(shared_ptr<meta_data_set>, string, int, int);

int main(int argc, char **argv)
{
    if (argc == 3) return ref_token_test_main(argc, argv);
    if (argc != 6) {
        fprintf(stderr, "usage: %s <matrix.mtx> <N> [<family> <sparse_coarsen> <coarsen>]\n", argv[0]);
        return 2;
    }
    const string mtx = argv[1], family = argv[3];
    const int s = atoi(argv[4]), c = atoi(argv[5]);
    set_config("DENSE_MATRIX_SIZE", atoi(argv[2])); // token_test.cc:1630
    shared_ptr<meta_data_set> ms = create_init_metadata_set_from_file(get_config()["DATA_SET"].as_string() + mtx, "");
    float t = 0;
    if (family == "thread_total") t = test_spmm_thread_total(ms, mtx, s, c);
    else if (family == "warp_total") t = test_spmm_warp_total(ms, mtx, s, c);
    else if (family == "warp_bitmap") t = test_spmm_warp_bitmap(ms, mtx, s, c);
    else if (family == "thread_bit_map") t = test_spmm_thread_bit_map(ms, mtx, s, c);
    else if (family == "warp_segment") t = test_spmm_warp_segment(ms, mtx, s, c);
    else if (family == "block_total") t = test_spmm_block_total(ms, mtx, s, c);
    else if (family == "block_bitmap") t = test_spmm_block_bitmap(ms, mtx, s, c);
    else {
        fprintf(stderr, "unknown family %s\n", family.c_str());
        return 2;
    }
    printf("%s : min time: %g\n", mtx.c_str(), t);
    return 0;
}
