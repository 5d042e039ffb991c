// t1_ht_dec.cu -- HTJ2K (ISO/IEC 15444-15) code-block decoder, sm_100a.  (under construction)
#include "t1.cuh"

int grkb_launch_t1_ht_decode(const T1Launch& l, cudaStream_t s)
{
	(void)l;
	(void)s;
	grkb_set_error("HT block decoder not built yet");
	return -1;
}
