/* empty: stands in for the UPMEM SDK header included by pim/PimManager.h:16 */
