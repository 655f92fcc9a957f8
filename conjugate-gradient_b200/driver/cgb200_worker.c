/* cgb200_worker -- one rank of a GPU team behind conj_grad (include/cgb200.h: cgb200_set_gpus,
 * CGB200_GPUS=N).  Started by libcgb200.so, never by users: it maps the row block the calling process
 * has put on this rank's GPU and runs the multi-rank loop with the others until told to quit. */
#include "cgb200.h"

int main(int argc, char **argv) { return cgb200_worker_main(argc, argv); }
