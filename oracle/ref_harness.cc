// oracle/ref_harness.cc -- TEST INFRASTRUCTURE ONLY.
//
// A small main of ours around the UNMODIFIED reference host code (linked from
// /root/reference by oracle/Makefile). It runs one of the reference driver's seven
// operator chains (token_test.cc:1003,1188,1253,1319,1393,1458,1518 -- the reference's own
// test_spmm_* functions, compiled with -Dmain=ref_token_test_main) on a .mtx file and
// dumps EVERY item of the resulting meta_data_set as text, one value per line, via the
// reference's own writer (meta_data_item::output_2_file, metadata_set.cc:193).
//
// The reference's execute_binary (executor.cc:6) would fork a compiler on the generated
// program, which cannot compile in this snapshot; this file
// supplies execute_binary instead (executor.cc is left out of the link) and uses the call
// as the hook "the whole chain has run" to dump the metadata set.
//
// usage: ref_harness <family> <matrix.mtx> <N> <sparse_coarsen> <coarsen> <dump_dir>
//   CWD must hold global_config.json, arm_code/ and data_source/ (tests/golden/gen_golden.py
//   prepares such a scratch dir).
#include "metadata_set.hpp"
#include "config.hpp"
#include "executor.hpp"
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include <sys/stat.h>

using namespace std;

float test_spmm_thread_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_bitmap(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_thread_bit_map(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_warp_segment(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_block_total(shared_ptr<meta_data_set>, string, int, int);
float test_spmm_block_bitmap(shared_ptr<meta_data_set>, string, int, int);

static shared_ptr<meta_data_set> g_set;
static string g_dump_dir;

// every key the seven chains can leave behind; meta_data_set has no public iterator, so probe
static const char *k_names[] = {
    "nz_row_indices", "nz_col_indices", "nz_vals", "origin_row_num", "origin_col_num", "origin_nnz_num",
    "begin_row_index", "end_row_index", "begin_col_index", "end_col_index", "original_nz_row_indices",
    "first_row_indices", "first_nz_indices", "first_row_indices_without_ending",
    "first_row_indices_relative_to_BMTB", "first_nz_indices_relative_to_BMTB",
    "first_row_indices_relative_to_BMW", "first_nz_indices_relative_to_BMW",
    "first_BMW_indices", "first_BMT_indices", "BMT_size_of_each_blk", "BMW_size_of_each_blk", "BMTB_size",
    "nz_row_indices_after_interlance_storage", "nz_col_indices_after_interlance_storage",
    "nz_vals_after_interlance_storage", "thread_bit_map", "bit_map_of_thread", "segment_ptr", "segment_offset",
    "segment_empty_flag", "segment_empty_row_indices", "first_row_indices_without_ending_relative_to_BMTB",
    "first_row_indices_without_ending_relative_to_BMW"};
static const POS_TYPE k_pos[] = {GLOBAL_META, TBLOCK_META, WARP_META, THREAD_META};

static void dump_all()
{
    mkdir(g_dump_dir.c_str(), 0755);
    FILE *idx = fopen((g_dump_dir + "/INDEX").c_str(), "w");
    for (POS_TYPE pos : k_pos)
        for (const char *name : k_names)
            for (int id = -1; id <= 0; id++)
            {
                if (!g_set->is_exist(pos, name, id))
                    continue;
                string key = get_metadata_item_name(pos, name, id);
                shared_ptr<meta_data_item> item = g_set->get_element(pos, name, id);
                item->output_2_file(g_dump_dir + "/" + key);
                shared_ptr<universal_array> arr = item->get_metadata_arr();
                // key, length, in-memory dtype, compressed dtype (code_source_data.cc:383)
                fprintf(idx, "%s %lu %s %s\n", key.c_str(), (unsigned long)arr->get_len(),
                        convert_data_type_to_string(arr->get_data_type()).c_str(),
                        convert_data_type_to_string(arr->get_compress_data_type()).c_str());
            }
    fclose(idx);
}

bool execute_binary(string execute_path, string filename, float &total_execute_time, float &gflops)
{
    // remember where the reference wrote its own hand-off (registered arrays + kernel_file.cpp)
    FILE *f = fopen((g_dump_dir + ".handoff").c_str(), "w");
    fprintf(f, "%s\n", execute_path.c_str());
    fclose(f);
    dump_all();
    total_execute_time = 0;
    gflops = 0;
    return true;
}

bool compile_spmv_code(string) { return true; }

int main(int argc, char **argv)
{
    if (argc != 7)
    {
        fprintf(stderr, "usage: %s <family> <matrix.mtx> <N> <S> <C> <dump_dir>\n", argv[0]);
        return 2;
    }
    string family = argv[1], mtx = argv[2];
    int n = atoi(argv[3]), s = atoi(argv[4]), c = atoi(argv[5]);
    g_dump_dir = argv[6];
    set_config("DENSE_MATRIX_SIZE", n); // token_test.cc:1630
    g_set = create_init_metadata_set_from_file(mtx, "");
    if (family == "load")
    {
        dump_all();
        return 0;
    }
    if (family == "thread_total") test_spmm_thread_total(g_set, mtx, s, c);
    else if (family == "warp_total") test_spmm_warp_total(g_set, mtx, s, c);
    else if (family == "warp_bitmap") test_spmm_warp_bitmap(g_set, mtx, s, c);
    else if (family == "thread_bit_map") test_spmm_thread_bit_map(g_set, mtx, s, c);
    else if (family == "warp_segment") test_spmm_warp_segment(g_set, mtx, s, c);
    else if (family == "block_total") test_spmm_block_total(g_set, mtx, s, c);
    else if (family == "block_bitmap") test_spmm_block_bitmap(g_set, mtx, s, c);
    else { fprintf(stderr, "unknown family %s\n", family.c_str()); return 2; }
    return 0;
}
