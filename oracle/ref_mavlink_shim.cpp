// oracle/ref_mavlink_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// extern "C" access to the reference's own (vendored, header-only) MAVLink C library under
// /root/reference/modules/mavlink_c: fills mavlink_optical_flow_rad_t exactly as src/mainloop.cpp:359-371 does and
// frames it exactly as src/mavlink_tcp.cpp:142-149 does.  Compiled by oracle/Makefile where the headers lie; no
// reference source is copied.  Output: oracle/_ref/libaof_ref_mavlink.so (git-ignored, travels to the GPU box).
#include <mavlink.h>

#include <cstdint>

extern "C" int ref_flow_rad_frame(uint64_t time_usec, int dt_us, float flow_x_ang, float flow_y_ang, double gx, double gy,
				  double gz, int flow_quality, uint8_t system_id, uint8_t component_id, uint8_t seq,
				  uint8_t *buffer /* MAVLINK_MAX_PACKET_LEN */)
{
	mavlink_optical_flow_rad_t msg;
	msg.time_usec = time_usec;
	msg.integration_time_us = dt_us;
	msg.integrated_x = flow_x_ang;
	msg.integrated_y = flow_y_ang;
	msg.integrated_xgyro = -gy; // switch to match pixel directions
	msg.integrated_ygyro = gx;  // switch to match pixel directions
	msg.integrated_zgyro = gz;
	msg.time_delta_distance_us = 0;
	msg.distance = -1.0;
	msg.temperature = 0;
	msg.sensor_id = 0;
	msg.quality = flow_quality;
	mavlink_get_channel_status(MAVLINK_COMM_0)->current_tx_seq = seq;
	mavlink_message_t m;
	mavlink_msg_optical_flow_rad_encode(system_id, component_id, &m, &msg);
	return (int)mavlink_msg_to_send_buffer(buffer, &m);
}
