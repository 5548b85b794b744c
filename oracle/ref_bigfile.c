/* TEST INFRASTRUCTURE ONLY.
 * Writes bigfile blocks and attributes with the reference's OWN vendored bigfile library (depends/bigfile/src/bigfile.c,
 * compiled unmodified where it lies by oracle/Makefile), the way petaio.cpp does (petaio.cpp:380-400 header attributes,
 * :590-700 blocks), so that the library's reader (shenqi_b200/csrc/sq_bigfile.cu) is pinned to files the
 * reference's writer produced. */
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
#include <depends/bigfile/src/bigfile.h>

static int open_or_create(BigFile * bf, const char * path)
{
    if(0 == big_file_open(bf, path)) return 0;
    return big_file_create(bf, path);
}

/* nrows x nmemb values of `dtype` (numpy style, e.g. "f8", "f4", "u8") split over nfile physical files */
int ref_bigfile_write_block(const char * path, const char * block, const char * dtype, int nmemb, const void * data, int64_t nrows, int nfile)
{
    BigFile bf; BigBlock bb; BigArray arr; BigBlockPtr ptr;
    if(open_or_create(&bf, path)) return -1;
    size_t * fsize = (size_t *) calloc(nfile > 0 ? nfile : 1, sizeof(size_t));
    for(int i = 0; i < nfile; i++)
        fsize[i] = (size_t) ((int64_t) (i + 1) * nrows / nfile - (int64_t) i * nrows / nfile);
    int rc = big_file_create_block(&bf, &bb, block, dtype, nmemb, nfile, fsize);
    free(fsize);
    if(rc) { big_file_close(&bf); return -2; }
    if(nrows > 0) {
        size_t dims[2] = {(size_t) nrows, (size_t) nmemb};
        big_array_init(&arr, (void *) data, dtype, 2, dims, NULL);
        if(big_block_seek(&bb, &ptr, 0) || big_block_write(&bb, &ptr, &arr)) { big_block_close(&bb); big_file_close(&bf); return -3; }
    }
    if(big_block_close(&bb)) { big_file_close(&bf); return -4; }
    return big_file_close(&bf);
}

/* an attribute on a block; the block is created empty (dtype NULL, as petaio's "Header") if it does not exist */
int ref_bigfile_set_attr(const char * path, const char * block, const char * name, const void * data, const char * dtype, int nmemb)
{
    BigFile bf; BigBlock bb;
    if(open_or_create(&bf, path)) return -1;
    if(big_file_open_block(&bf, &bb, block)) {
        if(big_file_create_block(&bf, &bb, block, NULL, 0, 0, NULL)) { big_file_close(&bf); return -2; }
    }
    if(big_block_set_attr(&bb, name, data, dtype, nmemb)) { big_block_close(&bb); big_file_close(&bf); return -3; }
    if(big_block_close(&bb)) { big_file_close(&bf); return -4; }
    return big_file_close(&bf);
}

const char * ref_bigfile_error(void) { return big_file_get_error_message(); }
